#include "vtk.h"
#include "../core/common.h"
#include "../host/xml.h"
#include <cstring>

namespace tclb {

static const char *B64 = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";

std::string base64_encode(const void *data, size_t len) {
    const unsigned char *p = (const unsigned char *)data;
    std::string out;
    out.reserve((len + 2) / 3 * 4);
    for (size_t i = 0; i < len; i += 3) {
        unsigned w = p[i] << 16;
        if (i + 1 < len) w |= p[i + 1] << 8;
        if (i + 2 < len) w |= p[i + 2];
        out += B64[(w >> 18) & 63];
        out += B64[(w >> 12) & 63];
        out += i + 1 < len ? B64[(w >> 6) & 63] : '=';
        out += i + 2 < len ? B64[w & 63] : '=';
    }
    return out;
}
bool base64_decode(const std::string &text, std::vector<unsigned char> &out) {
    static int T[256];
    static bool init = false;
    if (!init) { for (int i = 0; i < 256; i++) T[i] = -1; for (int i = 0; i < 64; i++) T[(unsigned char)B64[i]] = i; init = true; }
    out.clear();
    unsigned w = 0;
    int bits = 0;
    for (unsigned char c : text) {
        if (c == '=') { bits = 0; w = 0; continue; }       // padding ends a segment (length header + payload are separate)
        if (T[c] < 0) continue;
        w = (w << 6) | (unsigned)T[c];
        bits += 6;
        if (bits >= 8) { bits -= 8; out.push_back((unsigned char)((w >> bits) & 0xFF)); }
    }
    return true;
}

void VtkFileOut::open(const std::string &fn, bool index) {
    make_parent_dirs(fn);
    f_ = fopen(fn.c_str(), "w");
    if (!f_) fail("Could not open vtk file %s", fn.c_str());
    if (index) {
        std::string p = fn;
        size_t k = p.rfind(".vti");
        if (k != std::string::npos) p = p.substr(0, k);
        p += ".pvti";
        fp_ = fopen(p.c_str(), "w");
        if (!fp_) fail("Could not open (p)vtk file %s", p.c_str());
    }
}
void VtkFileOut::init(const Box &tot, const Box &r, const std::vector<std::pair<Box, std::string>> &pieces, const std::string &sel,
                      double spacing, double ox, double oy, double oz) {
    size_ = (size_t)r.size();
    fprintf(f_, "<?xml version=\"1.0\"?>\n<VTKFile type=\"ImageData\" version=\"0.1\" byte_order=\"LittleEndian\">\n");
    fprintf(f_, "<ImageData WholeExtent=\"%d %d %d %d %d %d\" Origin=\"%lg %lg %lg\" Spacing=\"%lg %lg %lg\">\n", r.x0, r.x0 + r.nx, r.y0,
            r.y0 + r.ny, r.z0, r.z0 + r.nz, ox, oy, oz, spacing, spacing, spacing);
    fprintf(f_, "<Piece Extent=\"%d %d %d %d %d %d\">\n", r.x0, r.x0 + r.nx, r.y0, r.y0 + r.ny, r.z0, r.z0 + r.nz);
    fprintf(f_, "<CellData %s>\n", sel.c_str());
    if (fp_) {
        fprintf(fp_, "<?xml version=\"1.0\"?>\n<VTKFile type=\"PImageData\" version=\"0.1\" byte_order=\"LittleEndian\">\n");
        fprintf(fp_, "<PImageData WholeExtent=\"%d %d %d %d %d %d\" Origin=\"%lg %lg %lg\" Spacing=\"%lg %lg %lg\">\n", tot.x0, tot.x0 + tot.nx,
                tot.y0, tot.y0 + tot.ny, tot.z0, tot.z0 + tot.nz, ox, oy, oz, spacing, spacing, spacing);
        for (auto &p : pieces)
            fprintf(fp_, "<Piece Extent=\"%d %d %d %d %d %d\" Source=\"%s\"/>\n", p.first.x0, p.first.x0 + p.first.nx, p.first.y0,
                    p.first.y0 + p.first.ny, p.first.z0, p.first.z0 + p.first.nz, p.second.c_str());
        fprintf(fp_, "<PCellData %s>\n", sel.c_str());
    }
}
void VtkFileOut::write_field(const std::string &name, const void *data, int elem, const char *tp, int comp) {
    if (!f_) fail("vtk writer used before opening a file");
    int len = (int)(size_ * elem * comp);      // elem = size of one component
    fprintf(f_, "<DataArray type=\"%s\" Name=\"%s\" format=\"binary\" encoding=\"base64\" NumberOfComponents=\"%d\">\n", tp, name.c_str(), comp);
    std::string h = base64_encode(&len, sizeof(int)), d = base64_encode(data, (size_t)len);
    fwrite(h.data(), 1, h.size(), f_);
    fwrite(d.data(), 1, d.size(), f_);
    fprintf(f_, "\n</DataArray>\n");
    if (fp_) fprintf(fp_, "<PDataArray type=\"%s\" Name=\"%s\" format=\"binary\" encoding=\"base64\" NumberOfComponents=\"%d\"/>\n", tp, name.c_str(), comp);
}
void VtkFileOut::finish() {
    fprintf(f_, "</CellData>\n</Piece>\n</ImageData>\n</VTKFile>\n");
    if (fp_) fprintf(fp_, "</PCellData>\n</PImageData>\n</VTKFile>\n");
}
void VtkFileOut::close() {
    if (f_) fclose(f_);
    if (fp_) fclose(fp_);
    f_ = fp_ = nullptr;
}

void VtiFile::load(const std::string &filename) {
    XmlDocument d;
    d.load_file(filename);
    XmlNode *img = d.root()->child("ImageData");
    if (!img) fail("%s: not a VTK ImageData file", filename.c_str());
    sscanf(img->attr("WholeExtent").c_str(), "%d %d %d %d %d %d", extent, extent + 1, extent + 2, extent + 3, extent + 4, extent + 5);
    sscanf(img->attr("Spacing", "1 1 1").c_str(), "%lf %lf %lf", spacing, spacing + 1, spacing + 2);
    sscanf(img->attr("Origin", "0 0 0").c_str(), "%lf %lf %lf", origin, origin + 1, origin + 2);
    XmlNode *piece = img->child("Piece");
    if (!piece) fail("%s: no Piece", filename.c_str());
    if (piece->has_attr("Extent"))
        sscanf(piece->attr("Extent").c_str(), "%d %d %d %d %d %d", extent, extent + 1, extent + 2, extent + 3, extent + 4, extent + 5);
    for (const char *sect : {"CellData", "PointData"}) {
        XmlNode *cd = piece->child(sect);
        if (!cd) continue;
        for (auto *a : cd->children("DataArray")) {
            VtiArray arr;
            arr.name = a->attr("Name");
            arr.type = a->attr("type");
            arr.components = atoi(a->attr("NumberOfComponents", "1").c_str());
            std::vector<unsigned char> raw;
            base64_decode(a->text(), raw);
            if (raw.size() < 4) fail("%s: array %s is empty", filename.c_str(), arr.name.c_str());
            int len;
            memcpy(&len, raw.data(), 4);
            if ((size_t)len + 4 > raw.size()) fail("%s: array %s is truncated", filename.c_str(), arr.name.c_str());
            arr.bytes.assign(raw.begin() + 4, raw.begin() + 4 + len);
            arrays.push_back(std::move(arr));
        }
    }
}
const VtiArray *VtiFile::find(const std::string &name) const {
    for (auto &a : arrays) if (a.name == name) return &a;
    return nullptr;
}
}  // namespace tclb
