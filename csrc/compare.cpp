// compare: field-by-field difference of two .vti/.pvti outputs in units of the machine epsilon of the
// array type; optional periodic shift.  `compare a.pvti b.pvti [eps_multiple] [dx dy dz]`
// (same contract as the reference's src/compare.cpp:242-298; built on this framework's own VTI reader).
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include "core/common.h"
#include "host/xml.h"
#include "io/vtk.h"

using namespace tclb;

struct Grid { int ext[6]; std::vector<VtiArray> arrays; };

static std::string dir_of(const std::string &p) {
    size_t k = p.find_last_of('/');
    return k == std::string::npos ? "" : p.substr(0, k + 1);
}
// load a .vti directly, or assemble the pieces of a .pvti into one grid
static Grid load_any(const std::string &fn) {
    Grid g;
    if (ends_with(fn, ".vti")) {
        VtiFile f;
        f.load(fn);
        memcpy(g.ext, f.extent, sizeof(g.ext));
        g.arrays = f.arrays;
        return g;
    }
    XmlDocument d;
    d.load_file(fn);
    XmlNode *img = d.root()->child("PImageData");
    if (!img) fail("%s: not a PImageData file", fn.c_str());
    sscanf(img->attr("WholeExtent").c_str(), "%d %d %d %d %d %d", g.ext, g.ext + 1, g.ext + 2, g.ext + 3, g.ext + 4, g.ext + 5);
    const int nx = g.ext[1] - g.ext[0], ny = g.ext[3] - g.ext[2], nz = g.ext[5] - g.ext[4];
    bool first = true;
    for (auto *p : img->children("Piece")) {
        VtiFile f;
        f.load(dir_of(fn) + p->attr("Source"));
        if (first) {
            for (auto &a : f.arrays) {
                VtiArray b = a;
                size_t es = a.bytes.size() / ((size_t)(f.extent[1] - f.extent[0]) * (f.extent[3] - f.extent[2]) * (f.extent[5] - f.extent[4]));
                b.bytes.assign((size_t)nx * ny * nz * es, 0);
                g.arrays.push_back(b);
            }
            first = false;
        }
        const int px = f.extent[1] - f.extent[0], py = f.extent[3] - f.extent[2], pz = f.extent[5] - f.extent[4];
        for (size_t k = 0; k < f.arrays.size(); k++) {
            size_t es = f.arrays[k].bytes.size() / ((size_t)px * py * pz);
            for (int z = 0; z < pz; z++)
                for (int y = 0; y < py; y++)
                    memcpy(&g.arrays[k].bytes[(((size_t)(z + f.extent[4] - g.ext[4]) * ny + (y + f.extent[2] - g.ext[2])) * nx + (f.extent[0] - g.ext[0])) * es],
                           &f.arrays[k].bytes[((size_t)z * py + y) * px * es], (size_t)px * es);
        }
    }
    return g;
}
static double elem(const VtiArray &a, size_t i) {
    if (a.type == "Float64") return ((const double *)a.bytes.data())[i];
    if (a.type == "Float32") return ((const float *)a.bytes.data())[i];
    if (a.type == "UInt8") return a.bytes[i];
    if (a.type == "UInt16") return ((const uint16_t *)a.bytes.data())[i];
    if (a.type == "UInt32") return ((const uint32_t *)a.bytes.data())[i];
    if (a.type == "Int32") return ((const int32_t *)a.bytes.data())[i];
    fail("unsupported array type %s", a.type.c_str());
}

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: compare a.(p)vti b.(p)vti [eps_multiple=8] [dx dy dz]\n"); return 2; }
    log_init(0, L_WARNING);
    try {
        double mult = argc > 3 ? atof(argv[3]) : 8.0;
        int sh[3] = {argc > 4 ? atoi(argv[4]) : 0, argc > 5 ? atoi(argv[5]) : 0, argc > 6 ? atoi(argv[6]) : 0};
        Grid A = load_any(argv[1]), B = load_any(argv[2]);
        const int nx = A.ext[1] - A.ext[0], ny = A.ext[3] - A.ext[2], nz = A.ext[5] - A.ext[4];
        if (nx != B.ext[1] - B.ext[0] || ny != B.ext[3] - B.ext[2] || nz != B.ext[5] - B.ext[4]) { printf("Different sizes\n"); return 1; }
        int bad = 0;
        for (auto &a : A.arrays) {
            const VtiArray *b = nullptr;
            for (auto &x : B.arrays) if (x.name == a.name) b = &x;
            if (!b) { printf("%s: missing in second file\n", a.name.c_str()); bad++; continue; }
            if (a.type != b->type || a.components != b->components) { printf("%s: type/components differ\n", a.name.c_str()); bad++; continue; }
            double eps = a.type == "Float64" ? DBL_EPSILON : (a.type == "Float32" ? FLT_EPSILON : 0.0);
            double worst = 0;
            for (int z = 0; z < nz; z++)
                for (int y = 0; y < ny; y++)
                    for (int x = 0; x < nx; x++) {
                        size_t i = ((size_t)z * ny + y) * nx + x;
                        size_t j = ((size_t)((z + sh[2] + nz) % nz) * ny + (y + sh[1] + ny) % ny) * nx + (x + sh[0] + nx) % nx;
                        for (int c = 0; c < a.components; c++) worst = std::max(worst, std::fabs(elem(a, i * a.components + c) - elem(*b, j * a.components + c)));
                    }
            bool ok = worst <= mult * eps;
            printf("%-20s %-8s max abs diff %.6le (%.2lf eps) %s\n", a.name.c_str(), a.type.c_str(), worst, eps > 0 ? worst / eps : worst, ok ? "OK" : "FAILED");
            if (!ok) bad++;
        }
        return bad ? 1 : 0;
    } catch (const std::exception &e) { fprintf(stderr, "compare: %s\n", e.what()); return 2; }
}
