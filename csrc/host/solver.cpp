#include "solver.h"
#include "../core/adjoint.h"
#include "../io/h5min.h"
#include "../core/common.h"
#include "../handlers/handler.h"
#include "../io/vtk.h"
#include "sampler.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>

namespace tclb {

void NameSet::add_from_string(const std::string &s, char sep) {
    for (auto &t : split(s, sep)) {
        std::string v = trim(t);
        if (v.empty()) continue;
        if (v == "all") all = true;
        else names.insert(v);
    }
}

Solver::Solver(const SolverOptions &o) : opt(o) {}
Solver::~Solver() {
    hands.clear();
    adjoint.reset();
    sampler.reset();
    geometry.clear();
    domain.reset();
}

// ---------------------------------------------------------------- case loading ----
int xpath_modify(XmlDocument &doc, XmlNode *config, const std::vector<std::string> &args);   // xpath_cli.cpp

void Solver::load_case(const std::string &file, const std::vector<std::string> &xargs) {
    conffile = file;
    set_output("");
    config.load_file(file);
    post_parse(xargs);
}
void Solver::load_case_string(const std::string &xml, const std::string &name, const std::vector<std::string> &xargs) {
    conffile = name;
    set_output("");
    config.parse(xml);
    post_parse(xargs);
}
void Solver::set_output(const std::string &out) {
    outpath = out + basename_noext(conffile);
    logf(L_NOTICE, "Setting output path to: %s", outpath.c_str());
}

void Solver::post_parse(const std::vector<std::string> &xargs) {
    XmlNode *cfg = config.child("CLBConfig");
    if (!cfg) fail("in %s: No \"CLBConfig\" element", conffile.c_str());
    // deprecated <Params A="1" B-zone="2"/>  ->  <Param name= value= [zone=] [gauge=]/>  (reference src/main.cpp:262-293)
    auto found = xpath_select(cfg, "//Params");
    if (!found.empty()) {
        logf(L_WARNING_IMP, "%zu deprecated Params elements found. Changing them to Param", found.size());
        cfg->set_attr("permissive", "true");
        for (auto &r : found) {
            XmlNode *node = r.node, *after = r.node;
            std::string gauge = node->attr("gauge");
            for (auto &a : node->attrs) {
                if (a.first == "gauge") continue;
                std::string par = a.first, zone;
                size_t i = par.find('-');
                if (i != std::string::npos) { zone = par.substr(i + 1); par = par.substr(0, i); }
                auto p = std::make_unique<XmlNode>(XmlType::Element, "Param");
                p->set_attr("name", par);
                p->set_attr("value", a.second);
                if (!zone.empty()) p->set_attr("zone", zone);
                if (!gauge.empty()) p->set_attr("gauge", gauge);
                after = node->parent->insert_after(std::move(p), after);
            }
            node->parent->remove(node);
        }
    }
    if (!xargs.empty()) {
        int st = xpath_modify(config, cfg, xargs);
        if (st == -444) { read_units(); stop_requested = true; return; }
        if (st != 0) fail("XPATH modification failed");
    }
    auto comments = xpath_select(cfg, "//comment()");
    if (!comments.empty()) {
        logf(L_OUTPUT, "Discarding %zu comments", comments.size());
        for (auto &c : comments) c.node->parent->remove(c.node);
    }
    read_units();
    XmlNode *geom = cfg->child("Geometry");
    if (!geom) fail("in %s: No \"Geometry\" element", conffile.c_str());
    int nx = round_half_away(units.alt(geom->attr("nx"), 1));
    int ny = round_half_away(units.alt(geom->attr("ny"), 1));
    int nz = round_half_away(units.alt(geom->attr("nz"), 1));
    logf(L_NOTICE, "Mesh size in config file: %dx%dx%d", nx, ny, nz);
    set_size(nx, ny, nz);
    set_output("");
}

void Solver::read_units() {
    XmlNode *cfg = config.child("CLBConfig");
    XmlNode *set = cfg ? cfg->child("Units") : nullptr;
    if (!set) { logf(L_WARNING, "No \"Units\" element in config file"); return; }
    int i = 1;
    for (auto *p : set->children("Param")) {
        if (!p->has_attr("value")) fail("Value not provided in Param in Units");
        if (!p->has_attr("gauge")) fail("Gauge not provided in Param in Units");
        std::string name = p->has_attr("name") ? p->attr("name") : "unnamed" + std::to_string(i);
        units.set_unit(name, units.read_text(p->attr("value")) / units.read_text(p->attr("gauge")));
        i++;
    }
    units.make_gauge();
    if (!domain || is_root()) units.print_gauge();
}

void Solver::set_size(int nx, int ny, int nz) {
    total = Box();
    total.nx = nx; total.ny = ny; total.nz = nz;
    logf(L_OUTPUT, "Global lattice size: %dx%dx%d", nx, ny, nz);
    int xs = opt.x_block > 0 ? opt.x_block : (opt.backend == TCLB_BACKEND_CUDA ? 32 : 1);
    total.nx += xs - 1 - ((total.nx - 1) % xs);          // pad x like the reference (src/Solver.cpp.Rt:273)
    std::string lib = opt.model_path.empty() ? Model::find_library(opt.model, opt.precision) : opt.model_path;
    model.reset(Model::load(lib));
    std::vector<int> lr = opt.local_ranks;
    if (lr.empty()) for (int r = 0; r < opt.nranks; r++) lr.push_back(r);
    std::vector<int> dev = opt.devices;
    if (dev.empty()) {
        int count = 1;
        if (opt.backend == TCLB_BACKEND_CUDA) {
            TCLB_CUDA(cudaGetDeviceCount(&count));
            if (count < 1) fail("no CUDA device visible");
            XmlNode *cfg = config.child("CLBConfig");
            bool over = cfg && cfg->attr_bool("oversubscribe_gpu");
            if ((int)lr.size() > count && !over)
                logf(L_WARNING, "%zu subdomains share %d GPU(s) in this process (oversubscribed)", lr.size(), count);
        }
        for (size_t i = 0; i < lr.size(); i++) dev.push_back((int)(i % count));
    }
    domain = std::make_unique<Domain>(model.get(), opt.backend, total.nx, total.ny, total.nz, opt.nranks, lr, dev);
    domain->use_direct_halo = opt.direct_halo;
    domain->finalize_wiring();
    for (int r : domain->local) {
        const Box &b = domain->dec.boxes[r];
        logf(L_OUTPUT, "Local lattice size (rank %d): %dx%dx%d", r, b.nx, b.ny, b.nz);
    }
    // launch-shape table of every kernel (reference src/GetThreads.cpp:7-45 prints the same at start-up)
    if (opt.backend == TCLB_BACKEND_CUDA && model->d->kernel_info) {
        logf(L_NOTICE, "  Threads  |      Action");
        for (int s = 0; s < model->d->n_stages; s++)
            for (int g = 0; g < 4; g++) {
                int regs = 0, smem = 0, lmem = 0, maxthr = 0;
                if (model->d->kernel_info(s, g, &regs, &smem, &lmem, &maxthr) != 0) continue;
                static const char *gm[] = {"NoGlobals", "IntegrateGlobals", "OnlyObjective", "Globals+Objective"};
                const int lnx = domain->lat[domain->local[0]]->nx, bx = lnx >= 128 ? 128 : (lnx >= 64 ? 64 : 32);
                logf(L_NOTICE, "  %3dx%-3d  | %s<%s>  regs %d, shared %d B, local %d B, max %d threads", bx, 128 / bx,
                     model->d->stages[s].name, gm[g], regs, smem, lmem, maxthr);
            }
    }
    // defaults (unit-aware strings)
    for (int i = 0; i < model->d->n_zonesettings; i++) domain->set_zone_setting(i, -1, units.alt(model->d->zonesettings[i].default_value));
    for (int i = 0; i < model->d->n_settings; i++)
        if (model->d->settings[i].derived_target < 0) domain->set_setting(i, units.alt(model->d->settings[i].default_value));
    for (int i = 0; i < model->d->n_settings; i++)
        if (model->d->settings[i].derived_target >= 0) domain->set_setting(i, units.alt(model->d->settings[i].default_value));
    geometry.clear();
    geometry.resize(opt.nranks);
    for (int r : domain->local) geometry[r] = std::make_unique<Geometry>(model.get(), domain->dec.boxes[r], total, units);
    setting_zones.clear();
    setting_zones["DefaultZone"] = 0;
}

// ---------------------------------------------------------------- naming ----
std::string Solver::out_iter_file(const std::string &name, const std::string &suffix, int rank) const {
    std::string s = strf("%s_%s_P%02d_%08d%s", outpath.c_str(), name.c_str(), rank < 0 ? first_rank() : rank, iter, suffix.c_str());
    make_parent_dirs(s);
    return s;
}
std::string Solver::out_global_file(const std::string &name, const std::string &suffix, int rank) const {
    std::string s = strf("%s_%s_P%02d%s", outpath.c_str(), name.c_str(), rank < 0 ? first_rank() : rank, suffix.c_str());
    make_parent_dirs(s);
    return s;
}
std::string Solver::out_iter_collective_file(const std::string &name, const std::string &suffix) const {
    std::string s = strf("%s_%s_%08d%s", outpath.c_str(), name.c_str(), iter, suffix.c_str());
    make_parent_dirs(s);
    return s;
}

// ---------------------------------------------------------------- log ----
void Solver::init_log(const std::string &filename) {
    const tclb_model_t *d = model->d;
    log_scales.clear();
    for (int i = 0; i < d->n_settings; i++) log_scales.push_back(1 / units.alt(d->settings[i].unit));
    for (int i = 0; i < d->n_zonesettings; i++) log_scales.push_back(1 / units.alt(d->zonesettings[i].unit));
    for (int i = 0; i < d->n_globals; i++) log_scales.push_back(1 / units.alt(d->globals[i].unit));
    for (const char *u : {"m", "s", "kg"}) log_scales.push_back(1 / units.alt(u));
    if (!is_root()) return;
    FILE *f = fopen(filename.c_str(), "wt");
    if (!f) fail("cannot write %s", filename.c_str());
    fprintf(f, "\"Iteration\",\"Time_si\",\"Walltime\",\"Optimization\"");
    for (int i = 0; i < d->n_settings; i++) fprintf(f, ",\"%s\",\"%s_si\"", d->settings[i].name, d->settings[i].name);
    for (int i = 0; i < d->n_zonesettings; i++)
        for (auto &z : setting_zones) fprintf(f, ",\"%s-%s\",\"%s-%s_si\"", d->zonesettings[i].name, z.first.c_str(), d->zonesettings[i].name, z.first.c_str());
    for (int i = 0; i < d->n_globals; i++) fprintf(f, ",\"%s\",\"%s_si\"", d->globals[i].name, d->globals[i].name);
    for (const char *s : {"dx", "dt", "dm"}) fprintf(f, ",\"%s_si\"", s);
    fprintf(f, "\n");
    fclose(f);
}
void Solver::write_log(const std::string &filename) {
    if (!is_root()) return;
    const tclb_model_t *d = model->d;
    Lattice *l = domain->lat[first_rank()];
    FILE *f = fopen(filename.c_str(), "at");
    if (!f) fail("cannot append to %s", filename.c_str());
    int j = 0;
    const int dt_idx = d->n_settings + d->n_zonesettings + d->n_globals + 1;
    fprintf(f, "%d, %.13le, %.13le, %d", iter, log_scales[dt_idx] * iter, walltime(), opt_iter);
    for (int i = 0; i < d->n_settings; i++, j++) { double v = l->settings[i]; fprintf(f, ", %.13le, %.13le", v, v * log_scales[j]); }
    for (int i = 0; i < d->n_zonesettings; i++, j++)
        for (auto &z : setting_zones) {
            double v = l->zs.get(i, z.second, l->zs.len() > 1 ? l->iter % l->zs.len() : 0);
            fprintf(f, ", %.13le, %.13le", v, v * log_scales[j]);
        }
    for (int i = 0; i < d->n_globals; i++, j++) { double v = domain->globals[i]; fprintf(f, ", %.13le, %.13le", v, v * log_scales[j]); }
    for (int i = 0; i < 3; i++, j++) fprintf(f, ", %.13le", log_scales[j]);
    fprintf(f, "\n");
    fclose(f);
}

// ---------------------------------------------------------------- writers ----
int Solver::write_vtk(const std::string &name, const NameSet &what, const Box &region_out) {
    logf(L_OUTPUT, "%8d it writing vtk", iter);
    const tclb_model_t *d = model->d;
    const char *rtype = model->real_size() == 8 ? "Float64" : "Float32";
    std::vector<std::pair<Box, std::string>> pieces;
    for (int r = 0; r < opt.nranks; r++) {
        Box b = domain->dec.boxes[r].intersect(region_out);
        std::string fn = out_iter_file(name, ".vti", r);
        size_t k = fn.find_last_of('/');
        pieces.push_back({b, k == std::string::npos ? fn : fn.substr(k + 1)});
    }
    double spacing = 1 / units.alt("m");
    for (int r : domain->local) {
        Lattice *l = domain->lat[r];
        Box reg = domain->dec.boxes[r].intersect(region_out);
        VtkFileOut vtk;
        vtk.open(out_iter_file(name, ".vti", r), r == 0);
        vtk.init(region_out, reg, pieces, "Scalars=\"rho\" Vectors=\"velocity\"", spacing, px * spacing, py * spacing, pz * spacing);
        Box loc = reg;
        loc.x0 -= l->region.x0; loc.y0 -= l->region.y0; loc.z0 -= l->region.z0;
        size_t n = (size_t)reg.size();
        std::vector<uint32_t> fl(n);
        l->get_flags(loc, fl.data());
        if (what.explicitly_in("flag")) {
            if (model->flag_size() == 2) { std::vector<uint16_t> s(fl.begin(), fl.end()); vtk.write_field("flag", s.data(), 2, "UInt16", 1); }
            else vtk.write_field("flag", fl.data(), 4, "UInt32", 1);
        }
        std::vector<unsigned char> small(n);
        for (int g = 0; g < d->n_nodegroups; g++) {
            const tclb_nodegroup_t &G = d->nodegroups[g];
            if ((what.all && G.save) || what.explicitly_in(G.name)) {
                for (size_t i = 0; i < n; i++) small[i] = (unsigned char)((fl[i] & G.mask) >> G.shift);
                vtk.write_field(G.name, small.data(), 1, "UInt8", 1);
            }
        }
        for (int q = 0; q < d->n_quantities; q++) {
            if (!what.in(d->quantities[q].name)) continue;
            double v = units.alt(d->quantities[q].unit);
            int comp = d->quantities[q].vector ? 3 : 1;
            std::vector<char> tmp(n * comp * model->real_size());
            l->get_quantity(q, loc, tmp.data(), 1 / v);
            vtk.write_field(d->quantities[q].name, tmp.data(), model->real_size(), rtype, comp);
        }
        // adjoint of a field: <VTK what="...,<field>_adj"/> after an <Adjoint> element (name with [ ] removed)
        if (adjoint) {
            for (int f = 0; f < d->n_fields; f++) {
                std::string nm = d->fields[f].name;
                nm.erase(std::remove_if(nm.begin(), nm.end(), [](char c) { return c == '[' || c == ']' || c == ' '; }), nm.end());
                nm += "_adj";
                if (!what.explicitly_in(nm)) continue;
                const int rs = model->real_size();
                std::vector<char> full((size_t)l->nx * l->ny * l->nz * rs), tmp(n * rs);
                adjoint->get_field(f, full.data());
                for (int z = 0; z < loc.nz; z++)
                    for (int y = 0; y < loc.ny; y++)
                        memcpy(tmp.data() + (((size_t)z * loc.ny + y) * loc.nx) * rs,
                               full.data() + ((((size_t)(z + loc.z0)) * l->ny + (y + loc.y0)) * l->nx + loc.x0) * rs, (size_t)loc.nx * rs);
                vtk.write_field(nm, tmp.data(), rs, rtype, 1);
            }
        }
        vtk.finish();
        vtk.close();
    }
    return 0;
}

// <HDF5>: one file for the whole output region (reference src/hdf5Lattice.cpp:26-339: datasets [nz][ny][nx] or
// [nz][ny][nx][3], one per node-type group and quantity, values in SI units; .xmf = 3DCoRectMesh index)
int Solver::write_hdf5(const std::string &name, const NameSet &what, const Box &region_out, bool as_double, bool write_xdmf, bool point_data, int deflate) {
    logf(L_OUTPUT, "%8d it writing hdf5", iter);
    const tclb_model_t *d = model->d;
    // Every subdomain in this process: ONE file of the whole region (what the reference's collective write produces,
    // src/hdf5Lattice.cpp:154-195).  One process per GPU: every process writes the pieces of its subdomains
    // (<name>_P<rank>_<iter>.h5) and the process that owns rank 0 writes an XDMF spatial collection over all pieces.
    bool all_local = true;
    for (int r = 0; r < opt.nranks; r++) if (!domain->lat[r]) all_local = false;
    struct PieceOut { Box box; std::string file; };
    std::vector<PieceOut> pieces;              // what the XDMF index describes
    std::string xattr_of_piece;                // attribute block with a %s for the file name (same for every piece up to the dims)
    auto base_of = [](const std::string &file) { return file.substr(file.find_last_of('/') == std::string::npos ? 0 : file.find_last_of('/') + 1); };

    auto write_region = [&](const Box &out, const std::vector<int> &ranks, const std::string &file, std::string &xattr) {
        const size_t N = (size_t)out.size();
        H5MinWriter h5;
        const std::string base = base_of(file);
        const std::string dims3 = strf("%d %d %d", out.nz, out.ny, out.nx);
        auto xdmf_attr = [&](const char *fname, int comp, int prec, const char *type) {
            xattr += strf("      <Attribute Center=\"%s\" %sName=\"%s\">\n        <DataItem DataType=\"%s\" Dimensions=\"%s%s\" Format=\"HDF\" Precision=\"%d\">%s:/%s</DataItem>\n      </Attribute>\n",
                          point_data ? "Node" : "Cell", comp == 3 ? "AttributeType=\"Vector\" " : "", fname, type, dims3.c_str(), comp == 3 ? " 3" : "", prec, base.c_str(), fname);
        };
        auto scatter = [&](const Box &reg, auto &&put) {
            for (int k = 0; k < reg.nz; k++)
                for (int j = 0; j < reg.ny; j++)
                    for (int i = 0; i < reg.nx; i++) {
                        size_t src = ((size_t)k * reg.ny + j) * reg.nx + i;
                        size_t dst = ((size_t)(k + reg.z0 - out.z0) * out.ny + (j + reg.y0 - out.y0)) * out.nx + (i + reg.x0 - out.x0);
                        put(src, dst);
                    }
        };
        std::vector<uint32_t> flags(N);
        for (int r : ranks) {
            Lattice *l = domain->lat[r];
            Box reg = domain->dec.boxes[r].intersect(out);
            if (reg.empty()) continue;
            Box loc = reg; loc.x0 -= l->region.x0; loc.y0 -= l->region.y0; loc.z0 -= l->region.z0;
            std::vector<uint32_t> fl((size_t)reg.size());
            l->get_flags(loc, fl.data());
            scatter(reg, [&](size_t s_, size_t t) { flags[t] = fl[s_]; });
        }
        std::vector<unsigned char> small(N);
        for (int g = 0; g < d->n_nodegroups; g++) {
            const tclb_nodegroup_t &G = d->nodegroups[g];
            if (!((what.all && G.save) || what.explicitly_in(G.name))) continue;
            for (size_t i = 0; i < N; i++) small[i] = (unsigned char)((flags[i] & G.mask) >> G.shift);
            h5.add(G.name, H5MinWriter::U8, {(uint64_t)out.nz, (uint64_t)out.ny, (uint64_t)out.nx}, small.data(), deflate);
            xdmf_attr(G.name, 1, 1, "UChar");
        }
        for (int q = 0; q < d->n_quantities; q++) {
            if (!what.in(d->quantities[q].name)) continue;
            const double v = units.alt(d->quantities[q].unit);
            const int comp = d->quantities[q].vector ? 3 : 1;
            std::vector<double> full(N * comp);
            for (int r : ranks) {
                Lattice *l = domain->lat[r];
                Box reg = domain->dec.boxes[r].intersect(out);
                if (reg.empty()) continue;
                Box loc = reg; loc.x0 -= l->region.x0; loc.y0 -= l->region.y0; loc.z0 -= l->region.z0;
                std::vector<char> tmp((size_t)reg.size() * comp * model->real_size());
                l->get_quantity(q, loc, tmp.data(), 1 / v);
                scatter(reg, [&](size_t s_, size_t t) {
                    for (int c = 0; c < comp; c++)
                        full[t * comp + c] = model->real_size() == 8 ? ((const double *)tmp.data())[s_ * comp + c] : (double)((const float *)tmp.data())[s_ * comp + c];
                });
            }
            std::vector<uint64_t> dims = {(uint64_t)out.nz, (uint64_t)out.ny, (uint64_t)out.nx};
            if (comp == 3) dims.push_back(3);
            if (as_double) h5.add(d->quantities[q].name, H5MinWriter::F64, dims, full.data(), deflate);
            else { std::vector<float> f32(full.begin(), full.end()); h5.add(d->quantities[q].name, H5MinWriter::F32, dims, f32.data(), deflate); }
            xdmf_attr(d->quantities[q].name, comp, as_double ? 8 : 4, "Float");
        }
        h5.write(file);
    };

    std::vector<std::string> xattrs;
    if (all_local) {
        std::vector<int> ranks;
        for (int r = 0; r < opt.nranks; r++) ranks.push_back(r);
        const std::string file = out_iter_collective_file(name, ".h5");
        std::string xa;
        write_region(region_out, ranks, file, xa);
        pieces.push_back({region_out, file});
        xattrs.push_back(xa);
    } else {
        for (int r = 0; r < opt.nranks; r++) {
            Box reg = domain->dec.boxes[r].intersect(region_out);
            if (reg.empty()) continue;
            const std::string file = out_iter_file(name, ".h5", r);
            std::string xa;
            if (domain->lat[r]) write_region(reg, {r}, file, xa);
            pieces.push_back({reg, file});
            xattrs.push_back(xa);
        }
        // the attribute blocks of pieces written elsewhere: same text as a local piece with that piece's dims and file name
        std::string pattern, pat_dims, pat_file;
        for (size_t k = 0; k < pieces.size(); k++)
            if (!xattrs[k].empty()) { pattern = xattrs[k]; pat_dims = strf("%d %d %d", pieces[k].box.nz, pieces[k].box.ny, pieces[k].box.nx); pat_file = base_of(pieces[k].file); break; }
        auto replace_all = [](std::string t, const std::string &a, const std::string &b) {
            for (size_t at = 0; (at = t.find(a, at)) != std::string::npos; at += b.size()) t.replace(at, a.size(), b);
            return t;
        };
        for (size_t k = 0; k < pieces.size(); k++)
            if (xattrs[k].empty() && !pattern.empty())
                xattrs[k] = replace_all(replace_all(pattern, "Dimensions=\"" + pat_dims, "Dimensions=\"" + strf("%d %d %d", pieces[k].box.nz, pieces[k].box.ny, pieces[k].box.nx)),
                                        pat_file + ":/", base_of(pieces[k].file) + ":/");
    }
    if (write_xdmf && domain->lat[0]) {
        const double m = units.alt("1m"), sec = units.alt("1s"), shift = point_data ? 0.5 : 0.0;
        const std::string xf = out_iter_collective_file(name, ".xmf");
        FILE *f = fopen(xf.c_str(), "w");
        if (!f) fail("cannot write %s", xf.c_str());
        fprintf(f, "<?xml version=\"1.0\"?>\n<Xdmf xmlns:xi=\"http://www.w3.org/2001/XInclude\" Version=\"3.0\">\n  <Domain>\n");
        const bool coll = pieces.size() > 1 || !all_local;
        if (coll) fprintf(f, "    <Grid Name=\"Lattice\" GridType=\"Collection\" CollectionType=\"Spatial\">\n      <Time Value=\"%.15g\"/>\n", iter / sec);
        for (size_t k = 0; k < pieces.size(); k++) {
            const Box &b = pieces[k].box;
            if (coll) fprintf(f, "    <Grid Name=\"Lattice_P%02zu\">\n", k);
            else fprintf(f, "    <Grid Name=\"Lattice\">\n      <Time Value=\"%.15g\"/>\n", iter / sec);
            fprintf(f, "      <Geometry Origin=\"\" Type=\"ORIGIN_DXDYDZ\">\n        <DataItem DataType=\"Float\" Dimensions=\"3\" Format=\"XML\" Precision=\"8\">%.15g %.15g %.15g</DataItem>\n",
                    (pz + shift + b.z0) / m, (py + shift + b.y0) / m, (px + shift + b.x0) / m);
            fprintf(f, "        <DataItem DataType=\"Float\" Dimensions=\"3\" Format=\"XML\" Precision=\"8\">%.15g %.15g %.15g</DataItem>\n      </Geometry>\n", 1 / m, 1 / m, 1 / m);
            if (point_data) fprintf(f, "      <Topology Dimensions=\"%d %d %d\" Type=\"3DCoRectMesh\"/>\n", b.nz, b.ny, b.nx);
            else fprintf(f, "      <Topology Dimensions=\"%d %d %d\" Type=\"3DCoRectMesh\"/>\n", b.nz + 1, b.ny + 1, b.nx + 1);
            fputs(xattrs[k].c_str(), f);
            fprintf(f, "    </Grid>\n");
        }
        if (coll) fprintf(f, "    </Grid>\n");
        fprintf(f, "  </Domain>\n</Xdmf>\n");
        fclose(f);
    }
    return 0;
}

int Solver::write_txt(const std::string &name, const NameSet &what, bool gzip) {
    logf(L_OUTPUT, "%8d it writing txt", iter);
    const tclb_model_t *d = model->d;
    for (int r : domain->local) {
        Lattice *l = domain->lat[r];
        std::string base = out_iter_file(name, "", r);
        Box loc; loc.nx = l->nx; loc.ny = l->ny; loc.nz = l->nz;
        if (r == 0) {
            FILE *f = fopen((base + "_info.txt").c_str(), "w");
            if (!f) fail("Cannot open file: %s_info.txt", base.c_str());
            fprintf(f, "dx: %lg\ndt: %lg\ndm: %lg\ndT: %lg\nsize: %lld\nNX: %d\nNY: %d\nNZ: %d\n", 1 / units.alt("m"), 1 / units.alt("s"),
                    1 / units.alt("kg"), 1 / units.alt("K"), loc.size(), loc.nx, loc.ny, loc.nz);
            fclose(f);
        }
        for (int q = 0; q < d->n_quantities; q++) {
            if (!what.in(d->quantities[q].name)) continue;
            std::string fn = base + "_" + d->quantities[q].name + ".txt";
            FILE *f = gzip ? popen(("gzip > " + fn + ".gz").c_str(), "w") : fopen(fn.c_str(), "w");
            if (!f) fail("Cannot open file: %s", fn.c_str());
            int comp = d->quantities[q].vector ? 3 : 1;
            size_t n = (size_t)loc.size();
            std::vector<char> tmp(n * comp * model->real_size());
            l->get_quantity(q, loc, tmp.data(), 1 / units.alt(d->quantities[q].unit));
            for (size_t i = 0; i < n; i++) {
                for (int c = 0; c < comp; c++) {
                    if (c) fprintf(f, " ");
                    if (model->real_size() == 4) fprintf(f, "%.8g", ((float *)tmp.data())[i * comp + c]);
                    else fprintf(f, "%.16lg", ((double *)tmp.data())[i * comp + c]);
                }
                fprintf(f, ((i + 1) % loc.nx) == 0 ? "\n" : " ");
            }
            if (gzip) pclose(f); else fclose(f);
        }
    }
    return 0;
}

int Solver::write_bin(const std::string &name) {
    logf(L_OUTPUT, "%8d it writing bin", iter);
    const tclb_model_t *d = model->d;
    for (int r : domain->local) {
        Lattice *l = domain->lat[r];
        std::string base = out_iter_file(name, "", r);
        Box loc; loc.nx = l->nx; loc.ny = l->ny; loc.nz = l->nz;
        for (int q = 0; q < d->n_quantities; q++) {
            int comp = d->quantities[q].vector ? 3 : 1;
            std::vector<char> tmp((size_t)loc.size() * comp * model->real_size());
            l->get_quantity(q, loc, tmp.data(), 1.0);
            std::string fn = base + "." + d->quantities[q].name + ".bin";
            FILE *f = fopen(fn.c_str(), "w");
            if (!f) fail("Cannot open file: %s", fn.c_str());
            fwrite(tmp.data(), 1, tmp.size(), f);
            fclose(f);
        }
    }
    return 0;
}

void Solver::save_solution(const std::string &prefix, std::vector<std::string> *files) {
    for (int r : domain->local) {
        std::string fn = strf("%s_%d.pri", prefix.c_str(), r);
        domain->lat[r]->save_pri(fn);
        if (files) files->push_back(fn);
    }
}
void Solver::load_solution(const std::string &prefix) {
    for (int r : domain->local) domain->lat[r]->load_pri(strf("%s_%d.pri", prefix.c_str(), r));
    // ghost layers come with the image; nothing else to refresh
}
void Solver::save_comp(const std::string &prefix, const std::string &comp) {
    // one parameter density as raw real_t[nx*ny*nz] per rank (reference src/Solver.cpp.Rt:494-523)
    int di = model->find_density(comp);
    if (di < 0) fail("Unknown component %s", comp.c_str());
    int fi = model->d->densities[di].field_index;
    for (int r : domain->local) {
        Lattice *l = domain->lat[r];
        std::vector<char> buf((size_t)l->nx * l->ny * l->nz * model->real_size());
        l->copy_field_out(fi, buf.data());
        std::string fn = strf("%s_%s_%d.comp", prefix.c_str(), comp.c_str(), r);
        make_parent_dirs(fn);
        FILE *f = fopen(fn.c_str(), "wb");
        if (!f) fail("cannot write %s", fn.c_str());
        fwrite(buf.data(), 1, buf.size(), f);
        fclose(f);
    }
}
void Solver::load_comp(const std::string &prefix, const std::string &comp) {
    int di = model->find_density(comp);
    if (di < 0) fail("Unknown component %s", comp.c_str());
    int fi = model->d->densities[di].field_index;
    for (int r : domain->local) {
        Lattice *l = domain->lat[r];
        std::vector<char> buf((size_t)l->nx * l->ny * l->nz * model->real_size());
        std::string fn = strf("%s_%s_%d.comp", prefix.c_str(), comp.c_str(), r);
        FILE *f = fopen(fn.c_str(), "rb");
        if (!f) fail("cannot read %s", fn.c_str());
        size_t n = fread(buf.data(), 1, buf.size(), f);
        fclose(f);
        if (n != buf.size()) fail("%s has the wrong size", fn.c_str());
        l->copy_field_in(fi, buf.data());
    }
}

// ---------------------------------------------------------------- parameters ----
void Solver::set_param(const std::string &par, const std::string &value, const std::string &zone, bool permissive, const std::string &gauge) {
    int zone_number = -1;
    if (!zone.empty()) {
        auto it = setting_zones.find(zone);
        if (it == setting_zones.end()) fail("Unknown zone %s (found while setting parameter %s)", zone.c_str(), par.c_str());
        zone_number = it->second;
    }
    double val = units.alt(value);
    if (par.empty()) {
        if (!gauge.empty()) { logf(L_OUTPUT, "Gauge without setting"); return; }
        fail("Setting name not specified in Param element");
    }
    int s = model->find_setting(par), z = model->find_zonesetting(par);
    if (s >= 0) {
        logf(L_OUTPUT, "Setting %s to %s (%lf)", par.c_str(), value.c_str(), val);
        domain->set_setting(s, val);
    } else if (z >= 0) {
        logf(L_OUTPUT, "Setting %s in zone %s (%d) to %s (%lf)", par.c_str(), zone.c_str(), zone_number, value.c_str(), val);
        domain->set_zone_setting(z, zone_number, val);
    } else if (permissive) {
        if (!gauge.empty()) logf(L_NOTICE, "Unknown setting %s with gauge", par.c_str());
        else logf(L_WARNING_IMP, "Unknown setting %s", par.c_str());
    } else fail("Unknown setting %s", par.c_str());
}

// ---------------------------------------------------------------- iteration ----
Adjoint *Solver::need_adjoint() {
    if (!adjoint) adjoint.reset(new Adjoint(domain.get()));
    return adjoint.get();
}
std::vector<int> Solver::parameter_fields() const {
    std::vector<int> out;
    for (int f = 0; f < model->d->n_fields; f++) if (model->d->fields[f].parameter) out.push_back(f);
    return out;
}
void Solver::iterate(int n, int type) {
    if (recording) {                   // inside <Adjoint type="unsteady">: Globals every iteration, inputs on the tape
        need_adjoint()->record(n);
        monitor();
        return;
    }
    bool last_glob = (type & ITER_LASTGLOB) != 0;
    if (type & ITER_INTEG) last_glob = false;
    if (sampler && sampler->active()) {
        // samples are taken after every iteration
        for (int i = 0; i < n; i++) {
            domain->iterate(1, last_glob && i == n - 1, false);
            sampler->update(domain.get());
        }
    } else domain->iterate(n, last_glob, (type & ITER_GLOBS) != 0);
    monitor();
}
void Solver::iterate_action(int action, int n, int type) {
    for (int i = 0; i < n; i++) {
        domain->run_action(action, TCLB_GLOBALS_NONE);
        domain->iter++;
        for (int r : domain->local) domain->lat[r]->iter = domain->iter;
    }
    monitor();
}
void Solver::monitor(bool force) {
    // throughput line in the reference's units (src/main.cpp:101-127): MLBUps over the padded global lattice
    double t = walltime();
    if (!force && t - mon_last_time_ < 1.0) return;
    domain->sync();
    t = walltime();
    int di = domain->iter - mon_last_iter_;
    if (di > 0 && mon_last_time_ > 0 && is_root()) {
        double lbups = (double)domain->nodes() * di / (t - mon_last_time_);
        logf(L_OUTPUT, "%8d it  %8.1f MLBUps  %7.2f GB/s", domain->iter, lbups / 1e6, lbups * model->bytes_per_update() / 1e9);
    }
    mon_last_time_ = t;
    mon_last_iter_ = domain->iter;
}
bool Solver::has_nan(const NameSet &what, const Box &region_chk) {
    bool bad = false;
    for (int q = 0; q < model->d->n_quantities && !bad; q++) {
        if (model->d->quantities[q].adjoint || !what.in(model->d->quantities[q].name)) continue;
        for (int r : domain->local) {
            Lattice *l = domain->lat[r];
            Box b = l->region.intersect(region_chk);
            if (b.empty()) continue;
            b.x0 -= l->region.x0; b.y0 -= l->region.y0; b.z0 -= l->region.z0;
            if (l->has_nan(q, b)) { bad = true; logf(L_NOTICE, "Checking %s discovered NaN", model->d->quantities[q].name); break; }
        }
    }
    if (domain->reduce_hook) { double v = bad ? 1 : 0; domain->reduce_hook(&v, 1, TCLB_OP_MAX); bad = v > 0; }
    return bad;
}
std::vector<double> Solver::get_quantity_global(int q, const Box &reg, double scale) {
    int comp = model->d->quantities[q].vector ? 3 : 1;
    std::vector<double> out((size_t)reg.size() * comp, 0.0);
    for (int r : domain->local) {
        Lattice *l = domain->lat[r];
        Box b = l->region.intersect(reg);
        if (b.empty()) continue;
        Box loc = b;
        loc.x0 -= l->region.x0; loc.y0 -= l->region.y0; loc.z0 -= l->region.z0;
        std::vector<char> tmp((size_t)b.size() * comp * model->real_size());
        l->get_quantity(q, loc, tmp.data(), scale);
        for (int z = 0; z < b.nz; z++)
            for (int y = 0; y < b.ny; y++)
                for (int x = 0; x < b.nx; x++)
                    for (int c = 0; c < comp; c++) {
                        size_t si = (((size_t)z * b.ny + y) * b.nx + x) * comp + c;
                        size_t di = (((size_t)(z + b.z0 - reg.z0) * reg.ny + (y + b.y0 - reg.y0)) * reg.nx + (x + b.x0 - reg.x0)) * comp + c;
                        out[di] = model->real_size() == 4 ? ((float *)tmp.data())[si] : ((double *)tmp.data())[si];
                    }
    }
    return out;
}

int Solver::run() {
    if (stop_requested) return 0;
    XmlNode *cfg = config.child("CLBConfig");
    HandlerPtr h = HandlerFactory::produce(cfg, this);
    if (!h) { logf(L_ERROR, "Something went wrong in xml run!"); return -1; }
    h.reset();
    domain->sync();
    if (integrator) { integrator->finish(); integrator.reset(); domain->particles_in_hook = nullptr; domain->particles_out_hook = nullptr; }
    if (debug_flags().profile) {
        // per-stage device time (TCLB_PROFILE=1): CUDA events around every launch of the first local subdomain
        Lattice *l = domain->lat[domain->local[0]];
        static const char *gm[] = {"", "+Globals", "+Objective", "+Globals+Objective"};
        logf(L_NOTICE_IMP, "%-32s %10s %12s %12s %10s", "stage", "launches", "total ms", "ms/launch", "GB/s");
        const double bytes = (double)l->nx * l->ny * l->nz * model->bytes_per_update();
        for (auto &kv : l->stage_stats) {
            const double per = kv.second.ms / std::max<long long>(1, kv.second.launches);
            logf(L_NOTICE_IMP, "%-32s %10lld %12.3f %12.5f %10.1f", (std::string(model->d->stages[kv.first.first].name) + gm[kv.first.second & 3]).c_str(),
                 kv.second.launches, kv.second.ms, per, per > 0 ? bytes / (per * 1e-3) / 1e9 : 0.0);
        }
    }
    if (is_root()) {
        double d = walltime();
        logf(L_OUTPUT, "Total duration: %lf s = %lf min = %lf h", d, d / 60, d / 3600);
    }
    return 0;
}
}  // namespace tclb
