// h5dump.cpp -- libhdf5-free writer of the reference's per-rank field dump (host code only, no CUDA).
//
// Replaces saveRankData (HDF5_output.h:11, HDF5_output.cpp:10-79): one file per rank per dump; for each owned tile a
// group /Tile_<globalID> with a 2-D dataset `fields` of shape (TY, TX) whose element is the 48-byte compound
// {Ex,Ey,Ez,Bx,By,Bz} (guards included) and three scalar int32 attributes tileRow, tileCol, currentRank.  The consumer
// contract is File_reader.py:82-117.  libhdf5 is not available in this image, so the file is assembled directly in the
// HDF5 1.x on-disk format that the reference's own files use (observed in Simulation/Fields/*.h5):
// superblock v0, old-style groups (v1 B-tree + local heap + symbol-table nodes, leaf K = 4, internal K = 16), v1 object
// headers, v1 attribute messages, dataspace v1, compound datatype v1, fill value v2, contiguous layout v3.
// Not byte-identical to a libhdf5-written file (allocation order differs, everything sits in one header chunk); the same
// objects, names, types and data bytes.  Checked by tests/h5mini.py (the parser that also reads the shipped dumps).
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <time.h>
#include <algorithm>
#include <string>
#include <vector>
#include "../../include/minipic_b200.h"

void mpic_set_error(const char* fmt, ...);

namespace {

const uint64_t UNDEF = ~0ull;
const int LEAF_K = 4, INTERNAL_K = 16;
const size_t SNOD_BYTES = 8 + 2 * LEAF_K * 40;
const size_t TREE_BYTES = 24 + (2 * INTERNAL_K + 1) * 8 + 2 * INTERNAL_K * 8;

struct Buf {
    std::vector<unsigned char> b;
    size_t size() const { return b.size(); }
    void pad8() { while (b.size() % 8) b.push_back(0); }
    size_t alloc(size_t n) { pad8(); size_t a = b.size(); b.resize(a + n, 0); return a; }
    void put(size_t at, const void* p, size_t n) { memcpy(&b[at], p, n); }
    void u8(size_t at, unsigned v) { b[at] = (unsigned char)v; }
    void u16(size_t at, unsigned v) { uint16_t x = (uint16_t)v; put(at, &x, 2); }
    void u32(size_t at, uint32_t v) { put(at, &v, 4); }
    void u64(size_t at, uint64_t v) { put(at, &v, 8); }
};

struct Msg { unsigned type, flags; std::vector<unsigned char> data; };

void app(std::vector<unsigned char>& d, const void* p, size_t n) { d.insert(d.end(), (const unsigned char*)p, (const unsigned char*)p + n); }
void app_u8(std::vector<unsigned char>& d, unsigned v) { d.push_back((unsigned char)v); }
void app_u16(std::vector<unsigned char>& d, unsigned v) { uint16_t x = (uint16_t)v; app(d, &x, 2); }
void app_u32(std::vector<unsigned char>& d, uint32_t v) { app(d, &v, 4); }
void app_u64(std::vector<unsigned char>& d, uint64_t v) { app(d, &v, 8); }
void app_pad8(std::vector<unsigned char>& d) { while (d.size() % 8) d.push_back(0); }

// v1 object header with all messages in one chunk; returns its address
size_t write_ohdr(Buf& f, const std::vector<Msg>& msgs) {
    size_t body = 0;
    for (const Msg& m : msgs) body += 8 + ((m.data.size() + 7) & ~(size_t)7);
    size_t at = f.alloc(16 + body);
    f.u8(at, 1); f.u16(at + 2, (unsigned)msgs.size()); f.u32(at + 4, 1); f.u32(at + 8, (uint32_t)body);
    size_t pos = at + 16;
    for (const Msg& m : msgs) {
        size_t sz = (m.data.size() + 7) & ~(size_t)7;
        f.u16(pos, m.type); f.u16(pos + 2, (unsigned)sz); f.u8(pos + 4, m.flags);
        if (!m.data.empty()) f.put(pos + 8, m.data.data(), m.data.size());
        pos += 8 + sz;
    }
    return at;
}

Msg attr_i32(const char* name, int32_t value) {
    Msg m; m.type = 0x000C; m.flags = 0;
    std::vector<unsigned char>& d = m.data;
    size_t nlen = strlen(name) + 1;
    app_u8(d, 1); app_u8(d, 0); app_u16(d, (unsigned)nlen); app_u16(d, 12); app_u16(d, 8);
    app(d, name, nlen); app_pad8(d);
    // datatype: fixed point v1, little-endian, signed, 4 bytes, bit offset 0, precision 32
    app_u8(d, 0x10); app_u8(d, 0x08); app_u8(d, 0); app_u8(d, 0); app_u32(d, 4); app_u16(d, 0); app_u16(d, 32); app_pad8(d);
    // dataspace v1 scalar
    app_u8(d, 1); app_u8(d, 0); app_u8(d, 0); for (int k = 0; k < 5; k++) app_u8(d, 0);
    app(d, &value, 4); app_pad8(d);
    return m;
}

void float64_type(std::vector<unsigned char>& d) {
    app_u8(d, 0x11); app_u8(d, 0x20); app_u8(d, 0x3f); app_u8(d, 0); app_u32(d, 8);
    app_u16(d, 0); app_u16(d, 64); app_u8(d, 52); app_u8(d, 11); app_u8(d, 0); app_u8(d, 52); app_u32(d, 1023);
}

Msg grid_compound_type() {
    static const char* names[6] = { "Ex", "Ey", "Ez", "Bx", "By", "Bz" };
    Msg m; m.type = 0x0003; m.flags = 1;
    std::vector<unsigned char>& d = m.data;
    app_u8(d, 0x16); app_u8(d, 6); app_u8(d, 0); app_u8(d, 0); app_u32(d, 48);
    for (int k = 0; k < 6; k++) {
        app(d, names[k], 3); app_pad8(d);                    // name, NUL-padded to 8
        app_u32(d, 8 * k);                                   // byte offset (HOFFSET(Grid, ...), HDF5_output.cpp:19-24)
        app_u8(d, 0); app_u8(d, 0); app_u8(d, 0); app_u8(d, 0);   // dimensionality 0 + reserved
        app_u32(d, 0); app_u32(d, 0);                        // permutation, reserved
        for (int q = 0; q < 4; q++) app_u32(d, 0);           // dimension sizes
        float64_type(d);
    }
    return m;
}

// local heap holding the given names (offset 0 is the empty string); fills name_off; returns heap header address
size_t write_heap(Buf& f, const std::vector<std::string>& names, std::vector<uint64_t>& name_off) {
    std::vector<unsigned char> seg(8, 0);
    name_off.clear();
    for (const std::string& n : names) {
        name_off.push_back(seg.size());
        app(seg, n.c_str(), n.size() + 1);
        app_pad8(seg);
    }
    size_t at = f.alloc(32 + seg.size());
    f.put(at, "HEAP", 4);
    f.u64(at + 8, seg.size());
    f.u64(at + 16, 1);                                       // H5HL_FREE_NULL: no free block
    f.u64(at + 24, at + 32);
    f.put(at + 32, seg.data(), seg.size());
    return at;
}

struct Sym { uint64_t name_off, ohdr; };

// group B-tree over name-sorted symbols: SNOD leaves of <= 2*LEAF_K entries under TREE nodes of <= 2*INTERNAL_K children
size_t write_group_btree(Buf& f, const std::vector<Sym>& syms) {
    struct Child { uint64_t addr, last_name; };
    std::vector<Child> level;
    for (size_t i = 0; i < syms.size() || (i == 0 && syms.empty()); i += 2 * LEAF_K) {
        size_t n = std::min<size_t>(2 * LEAF_K, syms.size() - i);
        size_t at = f.alloc(SNOD_BYTES);
        f.put(at, "SNOD", 4); f.u8(at + 4, 1); f.u16(at + 6, (unsigned)n);
        for (size_t k = 0; k < n; k++) {
            size_t e = at + 8 + 40 * k;
            f.u64(e, syms[i + k].name_off); f.u64(e + 8, syms[i + k].ohdr);    // cache type 0: nothing cached
        }
        level.push_back({ at, n ? syms[i + n - 1].name_off : 0 });
        if (syms.empty()) break;
    }
    for (int depth = 0;; depth++) {
        std::vector<Child> up;
        std::vector<size_t> nodes;
        for (size_t i = 0; i < level.size(); i += 2 * INTERNAL_K) {
            size_t n = std::min<size_t>(2 * INTERNAL_K, level.size() - i);
            size_t at = f.alloc(TREE_BYTES);
            f.put(at, "TREE", 4); f.u8(at + 4, 0); f.u8(at + 5, depth); f.u16(at + 6, (unsigned)n);
            f.u64(at + 8, UNDEF); f.u64(at + 16, UNDEF);
            // key[0] = smallest (the empty string for the leftmost node, else the previous node's last name)
            f.u64(at + 24, i == 0 ? 0 : level[i - 1].last_name);
            for (size_t k = 0; k < n; k++) {
                f.u64(at + 24 + 8 + 16 * k, level[i + k].addr);
                f.u64(at + 24 + 16 + 16 * k, level[i + k].last_name);
            }
            nodes.push_back(at);
            up.push_back({ at, level[i + n - 1].last_name });
        }
        for (size_t k = 0; k < nodes.size(); k++) {          // sibling links
            if (k > 0) f.u64(nodes[k] + 8, nodes[k - 1]);
            if (k + 1 < nodes.size()) f.u64(nodes[k] + 16, nodes[k + 1]);
        }
        if (up.size() == 1) return up[0].addr;
        level.swap(up);
    }
}

// old-style group object (heap + B-tree + header); `extra` = attribute messages
size_t write_group(Buf& f, std::vector<std::pair<std::string, uint64_t>> children, const std::vector<Msg>& extra,
                   uint64_t* btree_out = nullptr, uint64_t* heap_out = nullptr) {
    std::sort(children.begin(), children.end());             // B-tree order = strcmp order of the link names
    std::vector<std::string> names;
    for (auto& c : children) names.push_back(c.first);
    std::vector<uint64_t> off;
    size_t heap = write_heap(f, names, off);
    std::vector<Sym> syms;
    for (size_t k = 0; k < children.size(); k++) syms.push_back({ off[k], children[k].second });
    size_t btree = write_group_btree(f, syms);
    Msg st; st.type = 0x0011; st.flags = 0;
    app_u64(st.data, btree); app_u64(st.data, heap);
    std::vector<Msg> msgs; msgs.push_back(st);
    for (const Msg& m : extra) msgs.push_back(m);
    if (btree_out) *btree_out = btree;
    if (heap_out) *heap_out = heap;
    return write_ohdr(f, msgs);
}

}  // namespace

extern "C" int mpic_save_rank_data(const char* filename, int ntiles, const int* gids, const int* tile_rows,
                                   const int* tile_cols, const int* current_rank, int interior_nx, int interior_ny,
                                   int guard, const double* grid6) {
    if (!filename || ntiles < 0 || (ntiles && (!gids || !tile_rows || !tile_cols || !current_rank || !grid6)) ||
        interior_nx <= 0 || interior_ny <= 0 || guard < 0) {
        mpic_set_error("mpic_save_rank_data: bad argument");
        return MPIC_EINVAL;
    }
    const uint64_t TX = interior_nx + 2 * guard, TY = interior_ny + 2 * guard, patch = TX * TY * 48;
    Buf f;
    f.alloc(96);                                             // superblock + root symbol-table entry, filled at the end
    std::vector<std::pair<std::string, uint64_t>> root_children;
    const uint32_t now = (uint32_t)time(nullptr);
    for (int t = 0; t < ntiles; t++) {
        // raw data first (contiguous layout), then the dataset header, then the group
        size_t data = f.alloc(patch);
        f.put(data, grid6 + (size_t)t * TX * TY * 6, patch);
        std::vector<Msg> dm(5);
        dm[0].type = 0x0001; dm[0].flags = 0;                // dataspace v1, rank 2, max dims present ({totalY,totalX}, HDF5_output.cpp:35)
        app_u8(dm[0].data, 1); app_u8(dm[0].data, 2); app_u8(dm[0].data, 1); for (int k = 0; k < 5; k++) app_u8(dm[0].data, 0);
        app_u64(dm[0].data, TY); app_u64(dm[0].data, TX); app_u64(dm[0].data, TY); app_u64(dm[0].data, TX);
        dm[1] = grid_compound_type();
        dm[2].type = 0x0005; dm[2].flags = 1;                // fill value v2: alloc late, write if set, undefined
        { const unsigned char fv[8] = { 2, 2, 2, 1, 0, 0, 0, 0 }; app(dm[2].data, fv, 8); }
        dm[3].type = 0x0008; dm[3].flags = 0;                // layout v3 contiguous
        app_u8(dm[3].data, 3); app_u8(dm[3].data, 1); app_u64(dm[3].data, data); app_u64(dm[3].data, patch);
        dm[4].type = 0x0012; dm[4].flags = 0;                // modification time
        app_u8(dm[4].data, 1); app_u8(dm[4].data, 0); app_u8(dm[4].data, 0); app_u8(dm[4].data, 0); app_u32(dm[4].data, now);
        size_t dset = write_ohdr(f, dm);
        std::vector<Msg> attrs;
        attrs.push_back(attr_i32("tileRow", tile_rows[t]));
        attrs.push_back(attr_i32("tileCol", tile_cols[t]));
        attrs.push_back(attr_i32("currentRank", current_rank[t]));
        size_t grp = write_group(f, { { "fields", dset } }, attrs);
        root_children.push_back({ "Tile_" + std::to_string(gids[t]), grp });
    }
    uint64_t rbt = 0, rheap = 0;
    size_t root = write_group(f, root_children, {}, &rbt, &rheap);
    f.pad8();
    // superblock v0
    static const unsigned char sig[8] = { 0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n' };
    f.put(0, sig, 8);
    f.u8(13, 8); f.u8(14, 8);
    f.u16(16, LEAF_K); f.u16(18, INTERNAL_K);
    f.u64(24, 0); f.u64(32, UNDEF); f.u64(40, f.size()); f.u64(48, UNDEF);
    f.u64(56, 0); f.u64(64, root); f.u32(72, 1); f.u64(80, rbt); f.u64(88, rheap);
    FILE* fp = fopen(filename, "wb");
    if (!fp) { mpic_set_error("mpic_save_rank_data: cannot open %s", filename); return MPIC_EINVAL; }
    size_t w = fwrite(f.b.data(), 1, f.b.size(), fp);
    int rc = fclose(fp);
    if (w != f.b.size() || rc != 0) { mpic_set_error("mpic_save_rank_data: short write to %s", filename); return MPIC_EINVAL; }
    return 0;
}

// params.txt (PIC_2D.cpp:425-438): key=value lines in default ostream formatting (6 significant digits == %g)
extern "C" int mpic_write_params(const char* path, double box_x, double box_y, int nx, int ny, int guard, int interior_nx,
                                 int interior_ny, double sim_time, double dt, int total_steps) {
    FILE* fp = path ? fopen(path, "w") : nullptr;
    if (!fp) { mpic_set_error("mpic_write_params: cannot open %s", path ? path : "(null)"); return MPIC_EINVAL; }
    fprintf(fp, "box_x=%g\nbox_y=%g\nnx_global=%d\nny_global=%d\nguard=%d\ninterior_nx=%d\ninterior_ny=%d\nsim_time=%g\ndt=%g\ntotal_steps=%d\n",
            box_x, box_y, nx, ny, guard, interior_nx, interior_ny, sim_time, dt, total_steps);
    fclose(fp);
    return 0;
}
