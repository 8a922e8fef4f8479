// Minimal HDF5 writer (no libhdf5 in this image): version-0 superblock, one root group with a single symbol-table
// node, contiguous or chunked+deflate little-endian datasets of float32 / float64 / uint8.  Enough for what the <HDF5> element writes
// (one dataset per node-type group and per quantity) and readable by h5py / ParaView through the XDMF index.
// Reference: src/hdf5Lattice.cpp:26-339 (uses parallel libhdf5); chunking = slabs along the slowest dimension indexed by a version-1 B-tree, filter = deflate.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace tclb {

class H5MinWriter {
  public:
    enum Type { F32, F64, U8 };
    // data is copied; dims in C order (slowest first)
    // deflate 1..9: the dataset is stored chunked (slabs along the slowest dimension) and each chunk zlib-compressed;
    // 0 = contiguous
    void add(const std::string &name, Type t, const std::vector<uint64_t> &dims, const void *data, int deflate = 0);
    void write(const std::string &filename) const;          // throws (fail()) on I/O errors
    size_t datasets() const { return sets_.size(); }
    static constexpr int LEAF_K = 64;                       // symbol-table node capacity 2*LEAF_K entries
  private:
    struct Set { std::string name; Type type; std::vector<uint64_t> dims; std::vector<unsigned char> bytes; int deflate = 0; };
    std::vector<Set> sets_;
};

}  // namespace tclb
