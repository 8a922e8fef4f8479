// VTK ImageData writer: per-rank .vti (CellData, inline base64 with a 4-byte length header) and the
// .pvti index written by the owner of rank 0.  Same file format as the reference's vtkFileOut
// (src/vtkOutput.cpp:62-185, src/vtkLattice.cpp:9-56); the piece table is computed locally instead of
// broadcast, since every process knows the whole decomposition.
#pragma once
#include <cstdio>
#include <string>
#include <vector>
#include "../core/lattice.h"

namespace tclb {

std::string base64_encode(const void *data, size_t len);
bool base64_decode(const std::string &text, std::vector<unsigned char> &out);

class VtkFileOut {
  public:
    ~VtkFileOut() { close(); }
    // `pieces`: (region, source file basename) of every rank, written to the .pvti when write_index
    void open(const std::string &vti_filename, bool write_index);
    void init(const Box &total, const Box &piece, const std::vector<std::pair<Box, std::string>> &pieces,
              const std::string &selection, double spacing = 1.0, double ox = 0, double oy = 0, double oz = 0);
    void write_field(const std::string &name, const void *data, int elem_size, const char *type, int components);
    void finish();
    void close();
  private:
    FILE *f_ = nullptr, *fp_ = nullptr;
    size_t size_ = 0;
};

// reader used by `compare` and the tests
struct VtiArray { std::string name, type; int components = 1; std::vector<unsigned char> bytes; };
struct VtiFile {
    int extent[6] = {0, 0, 0, 0, 0, 0};
    double spacing[3] = {1, 1, 1}, origin[3] = {0, 0, 0};
    std::vector<VtiArray> arrays;
    void load(const std::string &filename);
    const VtiArray *find(const std::string &name) const;
};

}  // namespace tclb
