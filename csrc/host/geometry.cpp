#include "geometry.h"
#include "../core/common.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

namespace tclb {

int round_half_away(double v) { return v > 0 ? (int)(v + 0.5) : (int)(v - 0.5); }

const int Geometry::d3q27_vec[27][3] = {
    {0, 0, 0}, {1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {1, 1, 0}, {-1, 1, 0}, {0, -1, 0}, {1, -1, 0}, {-1, -1, 0},
    {0, 0, 1}, {1, 0, 1}, {-1, 0, 1}, {0, 1, 1}, {1, 1, 1}, {-1, 1, 1}, {0, -1, 1}, {1, -1, 1}, {-1, -1, 1},
    {0, 0, -1}, {1, 0, -1}, {-1, 0, -1}, {0, 1, -1}, {1, 1, -1}, {-1, 1, -1}, {0, -1, -1}, {1, -1, -1}, {-1, -1, -1}};

Geometry::Geometry(const Model *model, const Box &r, const Box &t, const UnitEnv &u) : region(r), total(t), model_(model), units_(u) {
    geom.assign((size_t)r.size(), 0u);
    zones["DefaultZone"] = 0;
}
void Geometry::activate_cuts() {
    if (Q.empty()) Q.assign((size_t)26 * region.size(), (uint16_t)65535);
}

static const char *kPredefined =
    "<Geometry>"
    "<Zone name='Inlet'><Box dx='0' dy='0' dz='0' fx='0' fy='-1' fz='-1'/></Zone>"
    "<Zone name='Outlet'><Box dx='-1' dy='0' dz='0' fx='-1' fy='-1' fz='-1'/></Zone>"
    "<Zone name='Channel'><Box dx='0' dy='0' dz='0' fx='-1' fy='0' fz='-1'/><Box dx='0' dy='-1' dz='0' fx='-1' fy='-1' fz='-1'/></Zone>"
    "<Zone name='Tunnel'><Box dx='0' dy='0' dz='0' fx='-1' fy='0' fz='-1'/><Box dx='0' dy='-1' dz='0' fx='-1' fy='-1' fz='-1'/>"
    "<Box dx='0' dy='0' dz='0' fx='-1' fy='-1' fz='0'/><Box dx='0' dy='0' dz='-1' fx='-1' fy='-1' fz='-1'/></Zone>"
    "</Geometry>";

XmlNode *Geometry::find_def(const char *element, const std::string &name) const {
    for (auto &k : root_->kids)
        if (k->type == XmlType::Element && k->name == element && k->attr("name") == name) return k.get();
    return nullptr;
}
int Geometry::ival(const XmlNode *n, const char *a, int def) const {
    if (!n->has_attr(a)) return def;
    return round_half_away(units_.alt(n->attr(a), def));
}
int Geometry::ival_prefixed(const std::string &text, char *prefix) const {
    std::string t = text;
    *prefix = t.empty() ? '+' : t[0];
    if (*prefix == '>' || *prefix == '<') t = t.substr(1);
    else *prefix = '+';
    return round_half_away(units_.alt(t));
}
void Geometry::set_flag(const std::string &name) {
    // built-in type table = the model's node types (+ None)
    if (XmlNode *t = find_def("Type", name)) {
        fg_ = (uint32_t)atol(t->attr("value", "0").c_str());
        set_mask(t->attr("mask"));
        mode_ = OVERWRITE;
        return;
    }
    const tclb_nodetype_t *nt = model_->find_nodetype(name);
    if (!nt) fail("Unknown flag (in xml): %s", name.c_str());
    fg_ = nt->value;
    set_mask(nt->group);
    mode_ = OVERWRITE;
}
void Geometry::set_mask(const std::string &name) {
    if (XmlNode *t = find_def("Mask", name)) { fg_mask_ = (uint32_t)atol(t->attr("value", "-1").c_str()); return; }
    if (name == "NONE") { fg_mask_ = 0; return; }
    const tclb_nodegroup_t *g = model_->find_nodegroup(name);
    if (!g) fail("Unknown mask (in xml): %s", name.c_str());
    fg_mask_ = g->mask;
}
void Geometry::set_mode(const std::string &m) {
    if (m == "overwrite") mode_ = OVERWRITE;
    else if (m == "fill") mode_ = FILL;
    else if (m == "change") mode_ = CHANGE;
    else fail("Unknown mode (in xml): %s", m.c_str());
}
void Geometry::set_zone(const std::string &name) {
    int id;
    auto it = zones.find(name);
    if (it != zones.end()) id = it->second;
    else { id = (int)zones.size(); zones[name] = id; }
    if (id >= model_->d->zone_max) fail("too many setting zones (%d); the model leaves room for %d", id + 1, model_->d->zone_max);
    const uint32_t zmask = (uint32_t)model_->d->zone_max << model_->d->zone_shift;
    fg_ = (fg_ & ~zmask) | ((uint32_t)id << model_->d->zone_shift);
    fg_mask_ |= zmask;
}

Box Geometry::get_region(const XmlNode *n) const {
    Box r;
    if (!n) { r.nx = r.ny = r.nz = 1; return r; }
    if (n->name == "Geometry") return total;
    r = get_region(n->parent);
    struct Ax { const char *d, *f, *n; int Box::*o; int Box::*s; };
    static const Ax axes[3] = {{"dx", "fx", "nx", &Box::x0, &Box::nx}, {"dy", "fy", "ny", &Box::y0, &Box::ny}, {"dz", "fz", "nz", &Box::z0, &Box::nz}};
    for (const Ax &a : axes) {
        if (n->has_attr(a.d)) {
            char side;
            int w = ival_prefixed(n->attr(a.d), &side);
            if (side == '<') w = r.*(a.s) + w;
            else if (side == '+' && w < 0) w = r.*(a.s) + w;
            r.*(a.o) += w;
            r.*(a.s) -= w;
        }
    }
    for (const Ax &a : axes) {
        if (n->has_attr(a.f)) {
            int w = round_half_away(units_.alt(n->attr(a.f)));
            if (w < 0) w = r.*(a.s) + w + r.*(a.o);
            r.*(a.s) = w - r.*(a.o) + 1;
        }
    }
    for (const Ax &a : axes)
        if (n->has_attr(a.n)) r.*(a.s) = round_half_away(units_.alt(n->attr(a.n)));
    return r;
}

static inline bool in_sphere(double x, double y, double z) {
    x = 2 * x - 1; y = 2 * y - 1; z = 2 * z - 1;
    return (x * x + y * y + z * z) < 1;
}
static inline bool in_wedge(double x, double y, double z, const std::string &type_in) {
    std::string type = type_in.empty() ? "UpperLeft" : type_in;
    bool xz = type.size() > 2 && type.compare(type.size() - 2, 2, "XZ") == 0;
    std::string base = xz ? type.substr(0, type.size() - 2) : type;
    double &v = xz ? z : y;
    if (base == "UpperLeft") {}
    else if (base == "UpperRight") x = 1. - x;
    else if (base == "LowerLeft") v = 1. - v;
    else if (base == "LowerRight") { x = 1. - x; v = 1. - v; }
    else return 0 < 1e-10;
    return (x - v) < 1e-10;
}

void Geometry::draw_zone(const std::string &name) {
    XmlNode *z = find_def("Zone", name);
    if (!z) fail("Unknown zone (in xml): %s", name.c_str());
    draw(z);
}

void Geometry::draw(XmlNode *node) {
    for (auto &kp : node->kids) {
        XmlNode *n = kp.get();
        if (n->type != XmlType::Element) continue;
        Box reg = get_region(n);
        const std::string &nm = n->name;
        auto U = [&](const char *a) { return units_.alt(n->attr(a)); };
        if (nm == "Box") {
            Box r = region.intersect(reg);
            for (int z = r.z0; z < r.z0 + r.nz; z++)
                for (int y = r.y0; y < r.y0 + r.ny; y++)
                    for (int x = r.x0; x < r.x0 + r.nx; x++) dot(x, y, z);
        } else if (nm == "HalfSphere" || nm == "Sphere") {
            bool half = nm == "HalfSphere";
            for (int x = reg.x0; x < reg.x0 + reg.nx; x++)
                for (int y = reg.y0; y < reg.y0 + reg.ny; y++)
                    for (int z = reg.z0; z < reg.z0 + reg.nz; z++) {
                        double fy = half ? 0.5 - (.5 + y - reg.y0) / reg.ny / 2. : (.5 + y - reg.y0) / reg.ny;
                        if (in_sphere((.5 + x - reg.x0) / reg.nx, fy, (.5 + z - reg.z0) / reg.nz)) dot(x, y, z);
                    }
        } else if (nm == "OffgridSphere") {
            double x0 = U("x"), y0 = U("y"), z0 = U("z"), Rx, Ry, Rz;
            if (!n->has_attr("R")) { Rx = U("Rx"); Ry = U("Ry"); Rz = U("Rz"); } else Rx = Ry = Rz = U("R");
            reg.x0 = (int)(x0 - Rx - 5); reg.y0 = (int)(y0 - Ry - 5); reg.z0 = (int)(z0 - Rz - 5);
            reg.nx = (int)(2 * Rx + 10); reg.ny = (int)(2 * Ry + 10); reg.nz = (int)(2 * Rz + 10);
            for (int x = reg.x0; x < reg.x0 + reg.nx; x++)
                for (int y = reg.y0; y < reg.y0 + reg.ny; y++)
                    for (int z = reg.z0; z < reg.z0 + reg.nz; z++) {
                        double xx = .5 + x - x0, yy = .5 + y - y0, zz = .5 + z - z0;
                        if (xx * xx / (Rx * Rx) + yy * yy / (Ry * Ry) + zz * zz / (Rz * Rz) < 1.) dot(x, y, z);
                    }
        } else if (nm == "OffgridPipe" || nm == "ZPipe") {
            double x0 = U("x"), y0 = U("y"), Rx, Ry;
            if (!n->has_attr("R")) { Rx = U("Rx"); Ry = U("Ry"); } else Rx = Ry = U("R");
            reg.x0 = (int)(x0 - Rx - 5); reg.y0 = (int)(y0 - Ry - 5);
            reg.nx = (int)(2 * Rx + 10); reg.ny = (int)(2 * Ry + 10);
            // the reference's ZPipe measures x from y0 (src/Geometry.cpp.Rt:960); kept for parity
            double xc = nm == "ZPipe" ? y0 : x0;
            for (int x = reg.x0; x < reg.x0 + reg.nx; x++)
                for (int y = reg.y0; y < reg.y0 + reg.ny; y++)
                    for (int z = reg.z0; z < reg.z0 + reg.nz; z++) {
                        double xx = .5 + x - xc, yy = .5 + y - y0;
                        if (xx * xx / (Rx * Rx) + yy * yy / (Ry * Ry) < 1.) dot(x, y, z);
                    }
        } else if (nm == "XPipe") {
            double y0 = U("y"), z0 = U("z"), Ry, Rz;
            if (!n->has_attr("R")) { Ry = U("Ry"); Rz = U("Rz"); } else Ry = Rz = U("R");
            reg.y0 = (int)(y0 - Ry - 5); reg.z0 = (int)(z0 - Rz - 5);
            reg.ny = (int)(2 * Ry + 10); reg.nz = (int)(2 * Rz + 10);
            for (int y = reg.y0; y < reg.y0 + reg.ny; y++)
                for (int z = reg.z0; z < reg.z0 + reg.nz; z++)
                    for (int x = reg.x0; x < reg.x0 + reg.nx; x++) {
                        double yy = .5 + y - y0, zz = .5 + z - z0;
                        if (zz * zz / (Rz * Rz) + yy * yy / (Ry * Ry) < 1.) dot(x, y, z);
                    }
        } else if (nm == "YPipe") {
            double y0 = U("y"), z0 = U("z"), Rx, Rz;
            if (!n->has_attr("R")) { Rx = U("Rx"); Rz = U("Rz"); } else Rx = Rz = U("R");
            double x0 = U("x");
            reg.x0 = (int)(x0 - Rx - 5); reg.z0 = (int)(z0 - Rz - 5);
            reg.nx = (int)(2 * Rx + 10); reg.nz = (int)(2 * Rz + 10);
            for (int y = reg.y0; y < reg.y0 + reg.ny; y++)
                for (int z = reg.z0; z < reg.z0 + reg.nz; z++)
                    for (int x = reg.x0; x < reg.x0 + reg.nx; x++) {
                        double xx = .5 + x - y0, zz = .5 + z - z0;     // (sic) reference uses y0 here
                        if (zz * zz / (Rz * Rz) + xx * xx / (Rx * Rx) < 1.) dot(x, y, z);
                    }
        } else if (nm == "XAnnulus") {
            double y0 = U("y"), z0 = U("z"), Ryo, Rzo, Ryi, Rzi;
            if (!n->has_attr("Ro")) { Ryi = U("Ry_i"); Rzi = U("Rz_i"); Ryo = U("Ry_o"); Rzo = U("Rz_o"); }
            else { Ryo = Rzo = U("Ro"); Ryi = Rzi = U("Ri"); }
            reg.y0 = (int)(y0 - Ryo - 5); reg.z0 = (int)(z0 - Rzo - 5);
            reg.ny = (int)(2 * Ryo + 10); reg.nz = (int)(2 * Rzo + 10);
            for (int y = reg.y0; y < reg.y0 + reg.ny; y++)
                for (int z = reg.z0; z < reg.z0 + reg.nz; z++)
                    for (int x = reg.x0; x < reg.x0 + reg.nx; x++) {
                        double yy = .5 + y - y0, zz = .5 + z - z0;
                        if ((zz * zz / (Rzo * Rzo) + yy * yy / (Ryo * Ryo) < 1.) && (zz * zz / (Rzi * Rzi) + yy * yy / (Ryi * Ryi) > 1.)) dot(x, y, z);
                    }
        } else if (nm == "Pipe") {
            for (int x = reg.x0; x < reg.x0 + reg.nx; x++)
                for (int y = reg.y0 - 1; y < reg.y0 + reg.ny + 1; y++)
                    for (int z = reg.z0 - 1; z < reg.z0 + reg.nz + 1; z++)
                        if (!in_sphere(0.5, (.5 + y - reg.y0) / reg.ny, (.5 + z - reg.z0) / reg.nz)) dot(x, y, z);
        } else if (nm == "PipeY") {
            for (int x = reg.x0 - 1; x < reg.x0 + reg.nx + 1; x++)
                for (int y = reg.y0; y < reg.y0 + reg.ny; y++)
                    for (int z = reg.z0 - 1; z < reg.z0 + reg.nz + 1; z++)
                        if (!in_sphere(0.5, (.5 + x - reg.x0) / reg.nx, (.5 + z - reg.z0) / reg.nz)) dot(x, y, z);
        } else if (nm == "PipeZ") {
            for (int z = reg.z0; z < reg.z0 + reg.nz; z++)
                for (int y = reg.y0 - 1; y < reg.y0 + reg.ny + 1; y++)
                    for (int x = reg.x0 - 1; x < reg.x0 + reg.nx + 1; x++)
                        if (!in_sphere(0.5, (.5 + y - reg.y0) / reg.ny, (.5 + x - reg.x0) / reg.nx)) dot(x, y, z);
        } else if (nm == "Cylinder") {
            for (int x = reg.x0; x < reg.x0 + reg.nx; x++)
                for (int y = reg.y0 - 1; y < reg.y0 + reg.ny + 1; y++)
                    for (int z = reg.z0 - 1; z < reg.z0 + reg.nz + 1; z++)
                        if (in_sphere((.5 + x - reg.x0) / reg.nx, (.5 + y - reg.y0) / reg.ny, 0.5)) dot(x, y, z);
        } else if (nm == "Wedge") {
            std::string type = n->attr("direction");
            for (int x = reg.x0; x < reg.x0 + reg.nx; x++)
                for (int y = reg.y0; y < reg.y0 + reg.ny; y++)
                    for (int z = reg.z0; z < reg.z0 + reg.nz; z++)
                        if (in_wedge((x - reg.x0) / (reg.nx - 1.), (y - reg.y0) / (reg.ny - 1.), (z - reg.z0) / (reg.nz - 1.), type)) dot(x, y, z);
        } else if (nm == "STL") {
            load_stl(reg, n);
        } else if (nm == "Sweep") {
            load_sweep(reg, n);
        } else if (nm == "Text") {
            Box crop = region.intersect(get_region(n->parent));
            if (!n->has_attr("file")) fail("No 'file' attribute in 'Text' element in xml conf");
            FILE *f = fopen(n->attr("file").c_str(), "rt");
            if (!f) fail("Could not open file: %s", n->attr("file").c_str());
            for (int x = reg.x0; x < reg.x0 + reg.nx; x++)
                for (int y = reg.y0; y < reg.y0 + reg.ny; y++)
                    for (int z = reg.z0; z < reg.z0 + reg.nz; z++) {
                        int v;
                        if (fscanf(f, "%d", &v) == EOF) { fclose(f); fail("File (%s) ended while reading", n->attr("file").c_str()); }
                        bool in = x >= crop.x0 && x < crop.x0 + crop.nx && y >= crop.y0 && y < crop.y0 + crop.ny && z >= crop.z0 && z < crop.z0 + crop.nz;
                        if (v != 0 && in) dot(x, y, z);
                    }
            fclose(f);
        } else {
            XmlNode *z = find_def("Zone", nm);
            if (!z) fail("Unknown geometry element: %s", nm.c_str());
            draw(z);
        }
    }
}

// ---- STL ------------------------------------------------------------------------------------
namespace {
#pragma pack(push, 1)
struct StlTri { float norm[3], p1[3], p2[3], p3[3]; short v; };
#pragma pack(pop)
static_assert(sizeof(StlTri) == 50, "binary STL triangle record");

bool solve4(double A[16], double b[4], double x[4]) {
    const int n = 4;
    for (int k = 0; k < n; k++) {
        int piv = k;
        for (int i = k + 1; i < n; i++) if (std::fabs(A[i * n + k]) > std::fabs(A[piv * n + k])) piv = i;
        if (std::fabs(A[piv * n + k]) < 1e-14) return false;
        if (piv != k) { for (int j = 0; j < n; j++) std::swap(A[k * n + j], A[piv * n + j]); std::swap(b[k], b[piv]); }
        for (int i = k + 1; i < n; i++) {
            double m = A[i * n + k] / A[k * n + k];
            for (int j = k; j < n; j++) A[i * n + j] -= m * A[k * n + j];
            b[i] -= m * b[k];
        }
    }
    for (int i = n - 1; i >= 0; i--) {
        double s = b[i];
        for (int j = i + 1; j < n; j++) s -= A[i * n + j] * x[j];
        x[i] = s / A[i * n + i];
    }
    return true;
}
// distance (0..1 -> 0..CUT_MAX) along direction d from node (x,y,z) to the triangle, NO_CUT if it misses:
// solve  w1 (P-p1) + w2 (P-p2) + w3 (P-p3) + t d = 0,  w1+w2+w3 = 1
uint16_t calc_cut(const StlTri &t, double x, double y, double z, double dx, double dy, double dz) {
    const double P[3] = {x, y, z}, D[3] = {dx, dy, dz};
    double A[16], b[4] = {0, 0, 0, 1}, r[4];
    for (int i = 0; i < 3; i++) {
        A[i * 4 + 0] = P[i] - t.p1[i]; A[i * 4 + 1] = P[i] - t.p2[i]; A[i * 4 + 2] = P[i] - t.p3[i]; A[i * 4 + 3] = D[i];
    }
    A[12] = A[13] = A[14] = 1; A[15] = 0;
    if (!solve4(A, b, r)) return 65535;
    if (r[0] < 0 || r[1] < 0 || r[2] < 0 || r[3] < 0 || r[3] > 1) return 65535;
    return (uint16_t)(r[3] * 1000);
}
}  // namespace

void Geometry::load_stl(const Box &reg, const XmlNode *n) {
    int inside_out = 0, axis = 1;
    if (!n->has_attr("file")) fail("No 'file' attribute in 'STL' element in xml conf");
    if (n->has_attr("side")) {
        std::string s = n->attr("side");
        if (s == "in") inside_out = 0; else if (s == "out") inside_out = 1; else if (s == "surface") inside_out = 2;
        else fail("'side' in 'STL' element have to be 'in', 'out' or 'surface'");
    }
    if (n->has_attr("ray_axis")) {
        if (inside_out > 1) fail("the 'ray_axis' setting in 'STL' makes sens only with 'in' or 'out' side");
        std::string s = n->attr("ray_axis");
        if (s == "x") axis = 0; else if (s == "y") axis = 1; else if (s == "z") axis = 2;
        else fail("'ray_axis' in 'STL' element have to be 'x', 'y' or 'z'");
    }
    FILE *f = fopen(n->attr("file").c_str(), "rb");
    if (!f) fail("'STL' element: %s doesn't exists or cannot be opened", n->attr("file").c_str());
    char header[80];
    int ntri = 0;
    if (fread(header, 80, 1, f) != 1 || fread(&ntri, sizeof(int), 1, f) != 1) { fclose(f); fail("STL file %s is truncated", n->attr("file").c_str()); }
    if (!strncmp(header, "solid", 5)) { fclose(f); fail("'STL' element %s is not in binary format!", n->attr("file").c_str()); }
    std::vector<StlTri> tri(ntri);
    size_t got = fread(tri.data(), sizeof(StlTri), ntri, f);
    fclose(f);
    if ((int)got != ntri) fail("STL file %s is truncated", n->attr("file").c_str());
    // transform: rotations about X,Y,Z, scale, translation, then snap to 1e-5 and shift off the grid lines
    for (int d = 0; d < 3; d++) {
        static const char *names[3] = {"Xrot", "Yrot", "Zrot"};
        if (!n->has_attr(names[d])) continue;
        double v = units_.alt(n->attr(names[d]));
        int i1 = (d + 1) % 3, i2 = (d + 2) % 3;
        for (auto &t : tri)
            for (float *p : {t.p1, t.p2, t.p3}) {
                double a = p[i1], b = p[i2];
                p[i1] = (float)(cos(v) * a - sin(v) * b);
                p[i2] = (float)(sin(v) * a + cos(v) * b);
            }
    }
    if (n->has_attr("scale")) {
        double v = units_.alt(n->attr("scale"));
        for (auto &t : tri) for (float *p : {t.p1, t.p2, t.p3}) for (int j = 0; j < 3; j++) p[j] = (float)(v * p[j]);
    }
    static const char *mv[3] = {"x", "y", "z"};
    for (int j = 0; j < 3; j++)
        if (n->has_attr(mv[j])) {
            double v = units_.alt(n->attr(mv[j]));
            for (auto &t : tri) for (float *p : {t.p1, t.p2, t.p3}) p[j] = (float)(p[j] + v);
        }
    const double sm[3] = {0.1403e-4, 0.1687e-4, 0.1987e-4};
    for (auto &t : tri) for (float *p : {t.p1, t.p2, t.p3}) for (int j = 0; j < 3; j++) {
        p[j] = (float)(round_half_away(p[j] * 1e5) * 1e-5);
        p[j] = (float)(p[j] + sm[j] - 0.5);
    }
    std::vector<char> lev;
    if (inside_out != 2) lev.assign((size_t)reg.size(), (char)inside_out);
    auto roff = [&](int x, int y, int z) { return ((size_t)(z - reg.z0) * reg.ny + (y - reg.y0)) * reg.nx + (x - reg.x0); };
    auto rin = [&](int x, int y, int z) { return x >= reg.x0 && x < reg.x0 + reg.nx && y >= reg.y0 && y < reg.y0 + reg.ny && z >= reg.z0 && z < reg.z0 + reg.nz; };
    size_t hits = 0, edge = 0, edge_bad = 0;
    for (const StlTri &t : tri) {
        int mn[3], mx[3];
        for (int j = 0; j < 3; j++) {
            mn[j] = (int)ceil(std::min({t.p1[j], t.p2[j], t.p3[j]})) - 1;
            mx[j] = (int)floor(std::max({t.p1[j], t.p2[j], t.p3[j]})) + 1;
        }
        if (inside_out == 2) {
            const size_t rs = (size_t)region.size();
            for (int x = mn[0]; x <= mx[0]; x++)
                for (int z = mn[2]; z <= mx[2]; z++)
                    for (int y = mn[1]; y <= mx[1]; y++) {
                        int lx = x - region.x0, ly = y - region.y0, lz = z - region.z0;
                        if (lx < 0 || ly < 0 || lz < 0 || lx >= region.nx || ly >= region.ny || lz >= region.nz) continue;
                        size_t k = ((size_t)lz * region.ny + ly) * region.nx + lx;
                        activate_cuts();
                        for (int d = 0; d < 26; d++) {
                            uint16_t q = calc_cut(t, x, y, z, d3q27_vec[d + 1][0], d3q27_vec[d + 1][1], d3q27_vec[d + 1][2]);
                            if (q < Q[rs * d + k]) Q[rs * d + k] = q;
                            if (q != 65535) dot(x, y, z);
                        }
                    }
        } else {
            const int a1 = axis, a2 = (axis + 1) % 3, a3 = (axis + 2) % 3;
            const int l1 = a1 == 0 ? reg.x0 : (a1 == 1 ? reg.y0 : reg.z0);
            double v1[2] = {t.p2[a2] - t.p1[a2], t.p2[a3] - t.p1[a3]}, v2[2] = {t.p3[a2] - t.p1[a2], t.p3[a3] - t.p1[a3]};
            double c0 = v1[0] * v2[1] - v1[1] * v2[0];
            const double dv[2] = {-0.5694552, 0.8220224};   // tie-break direction for rays through an edge
            double dc1 = (v1[0] * dv[1] - v1[1] * dv[0]) / c0, dc2 = (dv[0] * v2[1] - dv[1] * v2[0]) / c0, dc3 = -dc1 - dc2;
            for (int x2 = mn[a2]; x2 <= mx[a2]; x2++)
                for (int x3 = mn[a3]; x3 <= mx[a3]; x3++) {
                    double v[2] = {x2 - t.p1[a2], x3 - t.p1[a3]};
                    double c1 = (v1[0] * v[1] - v1[1] * v[0]) / c0, c2 = (v[0] * v2[1] - v[1] * v2[0]) / c0, c3 = 1. - c1 - c2;
                    int topo = 0;
                    const double cs[3] = {c1, c2, c3}, ds[3] = {dc1, dc2, dc3};
                    for (int q = 0; q < 3; q++) {
                        if (cs[q] == 0) { edge++; if (ds[q] > 0) topo++; else if (ds[q] == 0) edge_bad++; }
                        else if (cs[q] > 0) topo++;
                    }
                    if (topo != 3) continue;
                    hits++;
                    double h = t.p1[a1] * c3 + t.p2[a1] * c2 + t.p3[a1] * c1;
                    for (int x1 = l1; x1 <= h; x1++) {
                        int x, y, z;
                        if (axis == 0) { x = x1; y = x2; z = x3; }
                        else if (axis == 1) { x = x3; y = x1; z = x2; }
                        else { x = x2; y = x3; z = x1; }
                        if (rin(x, y, z)) lev[roff(x, y, z)]++;
                    }
                }
        }
    }
    if (inside_out != 2) {
        for (int x = reg.x0; x < reg.x0 + reg.nx; x++)
            for (int y = reg.y0; y < reg.y0 + reg.ny; y++)
                for (int z = reg.z0; z < reg.z0 + reg.nz; z++)
                    if (lev[roff(x, y, z)] % 2 == 1) dot(x, y, z);
        logf(L_OUTPUT, "STL: %d triangles, ray hits: %zu, edge hits: %zu (unresolved %zu)", ntri, hits, edge, edge_bad);
    }
}

// ---- Sweep (B-spline tube) ------------------------------------------------------------------
namespace {
double knot(int i, int n, int k) {
    if (i < k + 1) return 0;
    if (i < n) return 1.0 * (i - k) / (n - k);
    return 1;
}
double bspline(double x, std::vector<double> p, int k) {     // de Boor on a clamped uniform knot vector
    int n = (int)p.size();
    if (k > n - 1) k = n - 1;
    int i = (int)floor(x * (n - k)) + k;
    if (i > n - 1) i = n - 1;
    if (i < k) i = k;
    for (int j = k; j > 0; j--)
        for (int l = 0; l < j; l++) {
            double a = (x - knot(i - l, n, k)) / (knot(i - l + j, n, k) - knot(i - l, n, k));
            p[i - l] = a * p[i - l] + (1 - a) * p[i - l - 1];
        }
    return p[i];
}
}  // namespace

void Geometry::load_sweep(const Box &reg, const XmlNode *node) {
    int order = node->has_attr("order") ? atoi(node->attr("order").c_str()) : 1;
    double dl = 1e-4;
    if (node->has_attr("step")) dl = atof(node->attr("step").c_str());
    if (node->has_attr("steps")) dl = 1 / units_.alt(node->attr("steps"));
    double def_r = node->has_attr("r") ? units_.alt(node->attr("r")) : 1;
    std::vector<double> px, py, pz, pr;
    for (auto *p : node->children()) {
        if (p->name != "Point") { logf(L_NOTICE_IMP, "Unknown element %s in Sweep", p->name.c_str()); continue; }
        px.push_back(units_.alt(p->attr("x"))); py.push_back(units_.alt(p->attr("y"))); pz.push_back(units_.alt(p->attr("z")));
        pr.push_back(p->has_attr("r") ? units_.alt(p->attr("r"), def_r) : def_r);
    }
    int n = (int)px.size();
    if (n == 0) { logf(L_NOTICE_IMP, "No points in Sweep"); return; }
    if (order > n - 1) order = n - 1;
    double ox = -999, oy = -999, oz = -999, orr = -999;
    for (double l = 0; l < 1; l += dl) {
        double x0 = bspline(l, px, order), y0 = bspline(l, py, order), z0 = bspline(l, pz, order), r = bspline(l, pr, order);
        if (std::fabs(x0 - ox) < 0.25 && std::fabs(y0 - oy) < 0.25 && std::fabs(z0 - oz) < 0.25 && std::fabs(r - orr) < 0.25) continue;
        ox = x0; oy = y0; oz = z0; orr = r;
        Box b; b.x0 = (int)(x0 - r - 1); b.y0 = (int)(y0 - r - 1); b.z0 = (int)(z0 - r - 1);
        b.nx = b.ny = b.nz = (int)(2 * r + 2);
        Box t = reg.intersect(b);
        for (int x = t.x0; x < t.x0 + t.nx; x++)
            for (int y = t.y0; y < t.y0 + t.ny; y++)
                for (int z = t.z0; z < t.z0 + t.nz; z++)
                    if ((.5 + x - x0) * (.5 + x - x0) + (.5 + y - y0) * (.5 + y - y0) + (.5 + z - z0) * (.5 + z - z0) < r * r) dot(x, y, z);
    }
}

void Geometry::load(XmlNode *node) {
    logf(L_OUTPUT, "loading geometry ...");
    root_ = node;
    XmlDocument defs;
    defs.parse(kPredefined);
    for (auto *z : defs.root()->children()) {
        if (!z->has_attr("name")) continue;
        if (find_def(z->name.c_str(), z->attr("name"))) continue;
        node->prepend(z->clone());
    }
    for (size_t i = 0; i < node->kids.size(); i++) {
        XmlNode *n = node->kids[i].get();
        if (n->type != XmlType::Element) continue;
        if (n->name == "Zone" || n->name == "Type" || n->name == "Mask") continue;
        set_flag(n->name);
        for (auto &a : n->attrs) {
            if (a.first == "name") set_zone(a.second);
            else if (a.first == "mask") set_mask(a.second);
            else if (a.first == "mode") set_mode(a.second);
        }
        if (n->has_attr("zone")) draw_zone(n->attr("zone"));
        draw(n);
    }
}
}  // namespace tclb

#include "../io/vtk.h"
namespace tclb {
// <Geometry save="prefix">: dump flags (and cut distances) of a subdomain as <prefix>_P%02d.vti
// (reference src/Geometry.cpp.Rt:1202-1245)
void Geometry::write_vti(const std::string &prefix, int rank) const {
    VtkFileOut vtk;
    vtk.open(strf("%s_P%02d.vti", prefix.c_str(), rank), false);
    vtk.init(total, region, {}, "");
    if (model_->flag_size() == 2) {
        std::vector<uint16_t> s(geom.begin(), geom.end());
        vtk.write_field("geom", s.data(), 2, "UInt16", 1);
    } else vtk.write_field("geom", geom.data(), 4, "UInt32", 1);
    if (!Q.empty()) {
        const size_t rs = (size_t)region.size();
        std::vector<float> qn(rs);
        for (int d = 0; d < 26; d++) {
            for (size_t i = 0; i < rs; i++) qn[i] = (float)Q[rs * d + i] / 1000.f;
            vtk.write_field(strf("Qnorm%d%d%d", d3q27_vec[d + 1][0], d3q27_vec[d + 1][1], d3q27_vec[d + 1][2]), qn.data(), 4, "Float32", 1);
        }
        for (int d = 0; d < 26; d++)
            vtk.write_field(strf("Q%d%d%d", d3q27_vec[d + 1][0], d3q27_vec[d + 1][1], d3q27_vec[d + 1][2]), &Q[rs * d], 2, "UInt16", 1);
    }
    vtk.finish();
    vtk.close();
}
}  // namespace tclb
