// Geometry painter: turns the <Geometry> XML subtree into node-type flags (and wall-cut distances).
// Same input language as the reference's Geometry (src/Geometry.cpp.Rt:141-1245, predefined zones
// src/def.cpp.Rt:10-33): nested regions (dx,dy,dz,nx,ny,nz,fx,fy,fz; negative = from the far side,
// '<' '>' prefixes), brush = (flag, mask, mode overwrite/fill/change), named setting zones, primitives
// Box Sphere HalfSphere OffgridSphere OffgridPipe X/Y/ZPipe XAnnulus Pipe PipeY PipeZ Cylinder Wedge
// STL Sweep Text, and named Zone re-use.  Paints any sub-box of the global lattice, so every process
// only materialises the flags of the subdomains it owns.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>
#include "../core/lattice.h"
#include "../core/model.h"
#include "units.h"
#include "xml.h"

namespace tclb {

class Geometry {
  public:
    Geometry(const Model *model, const Box &region, const Box &total, const UnitEnv &units);
    std::vector<uint32_t> geom;            // [nz][ny][nx] of `region`
    std::vector<uint16_t> Q;               // optional cuts [26][nz][ny][nx]
    Box region, total;
    std::map<std::string, int> zones;      // setting zone name -> id ("DefaultZone" = 0)
    void load(XmlNode *geometry_node);     // throws on unknown elements/flags
    void write_vti(const std::string &prefix, int rank) const;
    static const int d3q27_vec[27][3];
  private:
    const Model *model_;
    UnitEnv units_;
    XmlNode *root_ = nullptr;
    uint32_t fg_ = 0, fg_mask_ = 0;
    enum Mode { OVERWRITE, FILL, CHANGE } mode_ = OVERWRITE;
    void set_flag(const std::string &name);
    void set_mask(const std::string &name);
    void set_mode(const std::string &name);
    void set_zone(const std::string &name);
    Box get_region(const XmlNode *n) const;
    void draw(XmlNode *n);
    void draw_zone(const std::string &name);
    void load_stl(const Box &reg, const XmlNode *n);
    void load_sweep(const Box &reg, const XmlNode *n);
    XmlNode *find_def(const char *element, const std::string &name) const;
    int ival(const XmlNode *n, const char *attr, int def) const;
    int ival_prefixed(const std::string &text, char *prefix) const;
    inline void dot(int x, int y, int z) {
        x -= region.x0; y -= region.y0; z -= region.z0;
        if (x < 0 || y < 0 || z < 0 || x >= region.nx || y >= region.ny || z >= region.nz) return;
        uint32_t &g = geom[((size_t)z * region.ny + y) * region.nx + x];
        if (g & fg_mask_) { if (mode_ == FILL) return; }
        else { if (mode_ == CHANGE) return; }
        g = (g & ~fg_mask_) | fg_;
    }
    void activate_cuts();
};

int round_half_away(double v);

}  // namespace tclb
