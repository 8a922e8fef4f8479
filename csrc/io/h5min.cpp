#include "h5min.h"
#include "../core/common.h"
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <zlib.h>

namespace tclb {
namespace {

const uint64_t UNDEF = ~0ull;

struct Buf {
    std::vector<unsigned char> b;
    void u8(unsigned v) { b.push_back((unsigned char)v); }
    void u16(unsigned v) { u8(v & 255); u8((v >> 8) & 255); }
    void u32(uint32_t v) { for (int i = 0; i < 4; i++) u8((v >> (8 * i)) & 255); }
    void u64(uint64_t v) { for (int i = 0; i < 8; i++) u8((unsigned)((v >> (8 * i)) & 255)); }
    void raw(const void *p, size_t n) { const unsigned char *c = (const unsigned char *)p; b.insert(b.end(), c, c + n); }
    void pad8() { while (b.size() % 8) u8(0); }
    void zeros(size_t n) { b.insert(b.end(), n, 0); }
    size_t size() const { return b.size(); }
};

size_t elem_size(H5MinWriter::Type t) { return t == H5MinWriter::F32 ? 4 : t == H5MinWriter::F64 ? 8 : 1; }

// one header message: type, data (padded to 8)
void message(Buf &o, unsigned type, const Buf &data) {
    size_t n = (data.size() + 7) & ~(size_t)7;
    o.u16(type); o.u16((unsigned)n); o.u8(0); o.u8(0); o.u8(0); o.u8(0);
    o.raw(data.b.data(), data.size());
    o.zeros(n - data.size());
}

Buf datatype_msg(H5MinWriter::Type t) {
    Buf d;
    if (t == H5MinWriter::U8) {
        d.u8(0x10);                         // version 1, class 0 (fixed point)
        d.u8(0); d.u8(0); d.u8(0);          // little endian, unsigned
        d.u32(1);
        d.u16(0); d.u16(8);                 // bit offset, precision
    } else {
        const bool dbl = t == H5MinWriter::F64;
        d.u8(0x11);                         // version 1, class 1 (floating point)
        d.u8(0x20);                         // little endian, mantissa normalisation: implied leading 1
        d.u8(dbl ? 63 : 31);                // sign bit position
        d.u8(0);
        d.u32(dbl ? 8 : 4);
        d.u16(0); d.u16(dbl ? 64 : 32);     // bit offset, precision
        d.u8(dbl ? 52 : 23); d.u8(dbl ? 11 : 8);   // exponent location, size
        d.u8(0); d.u8(dbl ? 52 : 23);       // mantissa location, size
        d.u32(dbl ? 1023 : 127);            // exponent bias
    }
    return d;
}

// chunk index constants: version-0 superblocks imply 32 for the "indexed storage internal node K"
const int CHUNK_K = 32;
size_t chunk_key_size(size_t rank) { return 8 + 8 * (rank + 1); }
size_t chunk_node_size(size_t rank) { return 24 + (2 * CHUNK_K + 1) * chunk_key_size(rank) + 2 * CHUNK_K * 8; }

// object header (version 1) of a dataset: contiguous (chunk0 == 0) or chunked along the slowest dimension with a deflate filter
Buf dataset_header(H5MinWriter::Type t, const std::vector<uint64_t> &dims, uint64_t data_addr, uint64_t data_bytes,
                   uint64_t chunk0 = 0, int deflate = 0) {
    Buf space;
    space.u8(1); space.u8((unsigned)dims.size()); space.u8(0); space.u8(0); space.u32(0);
    for (uint64_t d : dims) space.u64(d);
    Buf fill;
    if (chunk0) { fill.u8(2); fill.u8(3); fill.u8(2); fill.u8(0); }     // version 2, allocate incrementally (chunked), write if set, undefined
    else { fill.u8(2); fill.u8(1); fill.u8(2); fill.u8(0); }           // version 2, allocate early, write fill value if set, undefined
    Buf layout;
    if (chunk0) {
        layout.u8(3); layout.u8(2); layout.u8((unsigned)dims.size() + 1); layout.u64(data_addr);     // version 3, chunked, B-tree address
        layout.u32((uint32_t)chunk0);
        for (size_t i = 1; i < dims.size(); i++) layout.u32((uint32_t)dims[i]);
        layout.u32((uint32_t)elem_size(t));
    } else {
        layout.u8(3); layout.u8(1); layout.u64(data_addr); layout.u64(data_bytes);                   // version 3, contiguous
    }
    Buf msgs;
    message(msgs, 0x0001, space);
    message(msgs, 0x0003, datatype_msg(t));
    message(msgs, 0x0005, fill);
    if (chunk0) {
        Buf pipe;                                         // filter pipeline, version 1, one filter: deflate (id 1)
        pipe.u8(1); pipe.u8(1); pipe.zeros(6);
        pipe.u16(1); pipe.u16(0); pipe.u16(0); pipe.u16(1);   // id, name length 0, flags (mandatory), one client value
        pipe.u32((uint32_t)deflate); pipe.u32(0);             // the level, padding of the odd count of client values
        message(msgs, 0x000B, pipe);
    }
    message(msgs, 0x0008, layout);
    Buf h;
    h.u8(1); h.u8(0); h.u16(chunk0 ? 5 : 4); h.u32(1); h.u32((uint32_t)msgs.size()); h.u32(0);     // 16-byte prefix (incl. alignment pad)
    h.raw(msgs.b.data(), msgs.size());
    return h;
}

struct Chunked {
    uint64_t chunk0 = 0, nchunks = 0;                       // slab thickness along the slowest dimension, number of slabs
    std::vector<std::vector<unsigned char>> z;              // compressed slabs
    uint64_t payload_bytes = 0;
    int levels = 0;                                         // 1: one leaf node; 2: one internal node over several leaves
};

// index nodes first (root at the start of the payload), then the compressed slabs, each 8-aligned
Buf chunked_payload(const std::vector<uint64_t> &dims, size_t esize, const Chunked &c, uint64_t base) {
    const size_t rank = dims.size(), ksz = chunk_key_size(rank), nsz = chunk_node_size(rank);
    const uint64_t per = 2 * CHUNK_K;
    const uint64_t nleaf = (c.nchunks + per - 1) / per;
    const uint64_t nnodes = c.levels == 2 ? 1 + nleaf : 1;
    std::vector<uint64_t> addr(c.nchunks);
    uint64_t pos = base + nnodes * nsz;
    for (uint64_t i = 0; i < c.nchunks; i++) { pos = (pos + 7) & ~7ull; addr[i] = pos; pos += c.z[i].size(); }
    auto key = [&](Buf &o, uint64_t i) {                      // key of slab i; i == nchunks: the key past the last slab
        o.u32(i < c.nchunks ? (uint32_t)c.z[i].size() : 0); o.u32(0);
        o.u64(i * c.chunk0);
        for (size_t d = 1; d < rank; d++) o.u64(0);
        o.u64(0);                                               // offset in the element-size dimension
    };
    Buf out;
    auto leaf = [&](uint64_t first, uint64_t n, uint64_t left, uint64_t right) {
        const size_t start = out.size();
        out.raw("TREE", 4); out.u8(1); out.u8(0); out.u16((unsigned)n); out.u64(left); out.u64(right);
        for (uint64_t i = 0; i < n; i++) { key(out, first + i); out.u64(addr[first + i]); }
        key(out, first + n);
        out.zeros(nsz - (out.size() - start));
    };
    if (c.levels == 1) leaf(0, c.nchunks, UNDEF, UNDEF);
    else {
        const size_t start = out.size();
        out.raw("TREE", 4); out.u8(1); out.u8(1); out.u16((unsigned)nleaf); out.u64(UNDEF); out.u64(UNDEF);
        for (uint64_t l = 0; l < nleaf; l++) { key(out, l * per); out.u64(base + (1 + l) * nsz); }
        key(out, c.nchunks);
        out.zeros(nsz - (out.size() - start));
        for (uint64_t l = 0; l < nleaf; l++) {
            const uint64_t first = l * per, n = std::min<uint64_t>(per, c.nchunks - first);
            leaf(first, n, l > 0 ? base + l * nsz : UNDEF, l + 1 < nleaf ? base + (2 + l) * nsz : UNDEF);
        }
    }
    for (uint64_t i = 0; i < c.nchunks; i++) {
        out.zeros((size_t)(addr[i] - base) - out.size());
        out.raw(c.z[i].data(), c.z[i].size());
    }
    (void)ksz; (void)esize;
    return out;
}

Chunked make_chunks(const std::vector<uint64_t> &dims, size_t esize, const std::vector<unsigned char> &bytes, int level) {
    Chunked c;
    const uint64_t n0 = dims[0];
    uint64_t row = esize;
    for (size_t d = 1; d < dims.size(); d++) row *= dims[d];
    const uint64_t max_chunks = (uint64_t)(2 * CHUNK_K) * (2 * CHUNK_K);
    c.chunk0 = std::max<uint64_t>(1, (n0 + max_chunks - 1) / max_chunks);
    while (c.chunk0 * row > (1ull << 31) && c.chunk0 > 1) c.chunk0--;          // chunk sizes are 32-bit
    if (c.chunk0 * row >= (1ull << 32)) fail("H5MinWriter: one slab of %llu bytes is too large for a chunk", (unsigned long long)(c.chunk0 * row));
    c.nchunks = (n0 + c.chunk0 - 1) / c.chunk0;
    if (c.nchunks > max_chunks) fail("H5MinWriter: %llu chunks exceed the two-level index", (unsigned long long)c.nchunks);
    c.levels = c.nchunks <= (uint64_t)(2 * CHUNK_K) ? 1 : 2;
    std::vector<unsigned char> slab((size_t)(c.chunk0 * row));
    for (uint64_t i = 0; i < c.nchunks; i++) {
        const uint64_t first = i * c.chunk0, have = std::min<uint64_t>(c.chunk0, n0 - first);
        std::fill(slab.begin(), slab.end(), 0);                                // an edge slab is stored whole, padded
        memcpy(slab.data(), bytes.data() + first * row, (size_t)(have * row));
        uLongf zn = compressBound((uLong)slab.size());
        std::vector<unsigned char> z(zn);
        if (compress2(z.data(), &zn, slab.data(), (uLong)slab.size(), level) != Z_OK) fail("H5MinWriter: zlib failed");
        z.resize(zn);
        c.z.push_back(std::move(z));
    }
    const size_t nsz = chunk_node_size(dims.size());
    const uint64_t nleaf = (c.nchunks + 2 * CHUNK_K - 1) / (2 * CHUNK_K);
    uint64_t pos = (c.levels == 2 ? 1 + nleaf : 1) * nsz;
    for (auto &z : c.z) { pos = (pos + 7) & ~7ull; pos += z.size(); }
    c.payload_bytes = pos;
    return c;
}

}  // namespace

void H5MinWriter::add(const std::string &name, Type t, const std::vector<uint64_t> &dims, const void *data, int deflate) {
    if ((int)sets_.size() >= 2 * LEAF_K) fail("H5MinWriter: more than %d datasets in one file", 2 * LEAF_K);
    for (auto &s : sets_) if (s.name == name) fail("H5MinWriter: duplicate dataset %s", name.c_str());
    Set s;
    s.name = name; s.type = t; s.dims = dims;
    s.deflate = dims.empty() ? 0 : std::max(0, std::min(9, deflate));
    size_t n = elem_size(t);
    for (uint64_t d : dims) n *= (size_t)d;
    s.bytes.assign((const unsigned char *)data, (const unsigned char *)data + n);
    sets_.push_back(std::move(s));
}

void H5MinWriter::write(const std::string &filename) const {
    // symbol-table entries must be ordered by name
    std::vector<const Set *> order;
    for (auto &s : sets_) order.push_back(&s);
    std::sort(order.begin(), order.end(), [](const Set *a, const Set *b) { return a->name < b->name; });

    // ---- layout: superblock | root header | B-tree node | local heap header | heap data | SNOD | dataset headers | data
    const uint64_t a_super = 0, n_super = 96;
    const uint64_t a_root = a_super + n_super, n_root = 16 + 8 + 16;
    const uint64_t a_btree = a_root + n_root, n_btree = 24 + (2 * 16 + 1) * 8 + 2 * 16 * 8;
    const uint64_t a_heap = a_btree + n_btree, n_heaphdr = 32;
    // heap data: offset 0 = "" (8 zero bytes), then the names, then one free block
    Buf heap;
    heap.zeros(8);
    std::vector<uint64_t> name_off;
    for (auto *s : order) {
        name_off.push_back(heap.size());
        heap.raw(s->name.c_str(), s->name.size() + 1);
        heap.pad8();
    }
    const uint64_t free_off = heap.size();
    heap.u64(1);                 // no next free block
    heap.u64(32);                // size of this free block
    heap.zeros(16);
    const uint64_t a_heapdata = a_heap + n_heaphdr, n_heapdata = heap.size();
    const uint64_t a_snod = a_heapdata + n_heapdata, n_snod = 8 + (uint64_t)2 * LEAF_K * 40;
    uint64_t pos = a_snod + n_snod;
    std::vector<uint64_t> a_hdr, a_data;
    std::vector<Buf> hdrs;
    for (auto *s : order) { (void)s; a_hdr.push_back(0); a_data.push_back(0); }
    // headers first (need data addresses: compute sizes with a dummy pass)
    std::vector<uint64_t> hdr_size;
    std::vector<Chunked> chunked(order.size());
    for (size_t i = 0; i < order.size(); i++)
        if (order[i]->deflate > 0 && !order[i]->bytes.empty()) chunked[i] = make_chunks(order[i]->dims, elem_size(order[i]->type), order[i]->bytes, order[i]->deflate);
    auto payload_size = [&](size_t i) { return chunked[i].nchunks ? chunked[i].payload_bytes : (uint64_t)order[i]->bytes.size(); };
    for (size_t i = 0; i < order.size(); i++) hdr_size.push_back(dataset_header(order[i]->type, order[i]->dims, 0, 0, chunked[i].chunk0 * (chunked[i].nchunks ? 1 : 0), order[i]->deflate).size());
    for (size_t i = 0; i < order.size(); i++) { a_hdr[i] = pos; pos += hdr_size[i]; }
    for (size_t i = 0; i < order.size(); i++) { pos = (pos + 7) & ~7ull; a_data[i] = pos; pos += payload_size(i); }
    const uint64_t eof = pos;
    for (size_t i = 0; i < order.size(); i++)
        hdrs.push_back(dataset_header(order[i]->type, order[i]->dims, a_data[i], order[i]->bytes.size(), chunked[i].nchunks ? chunked[i].chunk0 : 0, order[i]->deflate));

    Buf f;
    // superblock, version 0
    const unsigned char sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
    f.raw(sig, 8);
    f.u8(0); f.u8(0); f.u8(0); f.u8(0); f.u8(0); f.u8(8); f.u8(8); f.u8(0);
    f.u16(LEAF_K); f.u16(16); f.u32(0);
    f.u64(0); f.u64(UNDEF); f.u64(eof); f.u64(UNDEF);
    f.u64(0); f.u64(a_root); f.u32(1); f.u32(0); f.u64(a_btree); f.u64(a_heap);       // root symbol-table entry
    // root group object header: one symbol-table message
    {
        Buf st; st.u64(a_btree); st.u64(a_heap);
        Buf msgs; message(msgs, 0x0011, st);
        f.u8(1); f.u8(0); f.u16(1); f.u32(1); f.u32((uint32_t)msgs.size()); f.u32(0);
        f.raw(msgs.b.data(), msgs.size());
    }
    // B-tree node (group node, leaf level, one child)
    {
        const size_t start = f.size();
        f.raw("TREE", 4); f.u8(0); f.u8(0); f.u16(order.empty() ? 0 : 1); f.u64(UNDEF); f.u64(UNDEF);
        f.u64(0);                                          // key 0: the empty name
        f.u64(a_snod);                                     // child 0
        f.u64(order.empty() ? 0 : name_off.back());        // key 1: largest name in child 0
        f.zeros(n_btree - (f.size() - start));
    }
    // local heap
    f.raw("HEAP", 4); f.u8(0); f.u8(0); f.u8(0); f.u8(0); f.u64(n_heapdata); f.u64(free_off); f.u64(a_heapdata);
    f.raw(heap.b.data(), heap.size());
    // symbol table node
    {
        const size_t start = f.size();
        f.raw("SNOD", 4); f.u8(1); f.u8(0); f.u16((unsigned)order.size());
        for (size_t i = 0; i < order.size(); i++) { f.u64(name_off[i]); f.u64(a_hdr[i]); f.u32(0); f.u32(0); f.zeros(16); }
        f.zeros(n_snod - (f.size() - start));
    }
    for (auto &h : hdrs) f.raw(h.b.data(), h.size());
    FILE *fp = fopen(filename.c_str(), "wb");
    if (!fp) fail("cannot write %s", filename.c_str());
    bool ok = fwrite(f.b.data(), 1, f.size(), fp) == f.size();
    uint64_t at = f.size();
    for (size_t i = 0; i < order.size() && ok; i++) {
        static const unsigned char z[8] = {0};
        if (a_data[i] > at) { ok = fwrite(z, 1, a_data[i] - at, fp) == a_data[i] - at; at = a_data[i]; }
        if (chunked[i].nchunks) {
            Buf pl = chunked_payload(order[i]->dims, elem_size(order[i]->type), chunked[i], a_data[i]);
            ok = ok && pl.size() == chunked[i].payload_bytes && fwrite(pl.b.data(), 1, pl.size(), fp) == pl.size();
            at += pl.size();
        } else {
            ok = ok && fwrite(order[i]->bytes.data(), 1, order[i]->bytes.size(), fp) == order[i]->bytes.size();
            at += order[i]->bytes.size();
        }
    }
    fclose(fp);
    if (!ok) fail("short write to %s", filename.c_str());
}

}  // namespace tclb
