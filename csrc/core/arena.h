// Bump allocator: every buffer of a subdomain is carved from ONE allocation (device or host),
// 256-byte aligned and zero-filled.  New implementation of the idea behind the reference's
// cudaPreAlloc / cudaAllocFinalize / cudaAllocFreeAll (src/cross.cu:38-117), sized for a
// 180 GB part: a 1024x512x512 d3q27 fp32 subdomain needs two 29 GB snapshots + flags in one slab,
// which also makes the whole subdomain exportable to peers with a single IPC handle.
#pragma once
#include <cstddef>
#include <string>
#include <vector>

namespace tclb {

class Arena {
  public:
    Arena(int backend, int device) : backend_(backend), device_(device), base_(nullptr), total_(0), finalized_(false) {}
    ~Arena();
    // register a request; returns slot id, pointer valid after finalize()
    int reserve(size_t bytes, const char *what);
    void finalize();
    void *ptr(int slot) const;
    size_t offset(int slot) const { return slots_[slot].off; }
    size_t size(int slot) const { return slots_[slot].bytes; }
    void *base() const { return base_; }
    size_t total() const { return total_; }
    int backend() const { return backend_; }
    int device() const { return device_; }
  private:
    struct Slot { size_t off, bytes; std::string what; };
    int backend_, device_;
    void *base_;
    size_t total_;
    bool finalized_;
    std::vector<Slot> slots_;
};

}  // namespace tclb
