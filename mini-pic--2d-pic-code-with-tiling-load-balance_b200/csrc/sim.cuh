// sim.cuh -- the handle behind mpic_sim* and the small macros shared by api.cu and ring.cu (internal).
#pragma once
#include "common.cuh"
#include <vector>

#define MPIC_PIPE_MAX 32
struct mpic_sim {
    mpic_params prm;
    FieldGeom G;
    PicConst P;
    YeeCoef cf;
    cudaStream_t st = nullptr;
    double* planes = nullptr;      // 6 field planes
    int64_t* jacc = nullptr;       // 3 current accumulators
    int64_t* rho = nullptr;        // 1 plane, diagnostics only (lazy)
    Particles pa[2] = {};
    int cur = 0;
    int64_t n = 0, cap = 0;
    int* cell_start = nullptr;     // n_entries + 1: first slot of every cell in the CURRENT sorted order
    int* cell_start_alt = nullptr; // n_entries + 1: the sort writes the new offsets here (the old ones are still being walked), then swaps
    int* cell_count = nullptr;     // n_entries: destination histogram of the last push; after the scan: arrival / scatter cursor
    int* mcnt = nullptr;           // 9 x ncells: per source cell and direction, particles moved by the last push (written by the push)
    int* offtab = nullptr;         // 9 x ncells: first slot of the particles moving (source cell, direction) in the new order
    unsigned char* dirs = nullptr; // cap: direction code of every stored particle's last move (written by the push)
    int* lv_off = nullptr;         // 2 x nx: first record of every slab-edge cell's leavers (nranks > 1)
    int* scan_tmp = nullptr;
    int n_entries = 0;             // ncells + 2 (leave-down, leave-up bins)
    CUtensorMap tmap;               // push kernel: tile + 1-cell ring box on `planes`
    bool has_tmap = false;
    // fused Yee step (yee.cu: fused_yee_kernel) ping-pongs between `planes` and `planes_alt`; after every fused step the two
    // blocks (and every map / peer pointer bound to them) are swapped, so `planes` always names the current fields
    double* planes_alt = nullptr;
    CUtensorMap tmap_alt, ftmap, ftmap_alt;   // push box on planes_alt; fused-step box on planes / planes_alt
    CUtensorMap tmap_pack, tmap_pack_alt;     // packed-walk push: part-row box on planes / planes_alt
    bool fused_ready = false;
    double* d_scratch = nullptr;   // 8 doubles
    int64_t tag_base = 0;
    int64_t launches = 0;
    int push_split = 4;
    int nsm = 148;
    bool counts_valid = false;     // cell_count holds the destination-cell histogram of the last push
    int* keys = nullptr;           // cap: sort key of stored particles in arbitrary order (uploads, arrivals); scratch of the segment sort
    int* perm = nullptr;           // cap: slot of the sorted order -> storage index in pa[cur]
    bool perm_valid = false;       // true: the sorted order is pa[cur] READ THROUGH perm (the sort moved no data; the next push
                                   // gathers); false: pa[cur][0..n) is itself in sorted order
    // ---- slab plumbing (nranks > 1)
    double* halo_send = nullptr;   // [2 sides][HALO_HDR + 3*g*pitch] values
    double* halo_recv = nullptr;
    int64_t halo_side = 0;         // values per side
    LeaveBufs lv = {};             // leaver records written by the push kernel
    double* arrive[2] = {};        // records received from the lower / upper neighbour
    int64_t arrived = 0;           // records absorbed into the tail [n, n+arrived) and not yet sorted in
    int64_t left = -1;             // leavers of the last push (host copy; -1: unknown)
    int64_t migrated_total = 0;    // records handed to the neighbours so far (harness counter)
    // ---- peer-memory ring (peer.cu)
    struct PeerSide {
        double* planes = nullptr;          // neighbour's 6 field planes
        double* planes_alt = nullptr;      // neighbour's other plane block (fused step ping-pong; swapped in lockstep)
        unsigned long long* jacc = nullptr; // neighbour's 3 current accumulators
        double* arrive = nullptr;          // neighbour's arrival block: [parity][from side][cap records]
        unsigned long long* ctl = nullptr; // neighbour's control block (flags + arrival counts)
        int64_t plane = 0;                 // neighbour's values per plane
        int nyl = 0;                       // neighbour's interior rows
        int lcap = 0;                      // neighbour's records per arrival buffer
        bool ipc = false;                  // pointers came from cudaIpcOpenMemHandle
    } peer[2];                             // 0 = lower neighbour, 1 = upper neighbour
    bool peer_on = false;
    double* arrive_blk = nullptr;          // own arrival block [2 parities][2 sides][lv.cap records] (arrive[] alias parity 0)
    unsigned long long* ctl = nullptr;     // own control block, written by the neighbours
    unsigned long long epoch[8] = {};      // per phase: last epoch signalled
    unsigned long long step_no = 0;        // PIC steps begun on the peer path (parity of the arrival buffers)
    long long* h_counts = nullptr;         // pinned + mapped: leave[2], arrive[2], error word (written by peer_sync_kernel)
    // MPIC_PEER_TRACE=1: per-phase device time of the peer step (events on the stream; adds one host wait per step)
    bool trace = false;
    cudaEvent_t tev[16] = {};
    double tacc[16] = {};
    long long tsteps = 0;
    int pending_ph = 3;
    bool b2_pending = false;               // B2 signalled, neighbours' B2 not awaited yet
    cudaEvent_t ev_counts = nullptr;
    cudaEvent_t ev_hand = nullptr;         // own hand-off finished (recorded on the side stream, awaited by the E sweep)
    // ---- pipelined host-buffer step (api.cu: pic_step_host_pipelined)
    cudaStream_t pipe_up = nullptr, pipe_down = nullptr;
    // single slab: the sort after a push runs on its own stream, next to the field update (neither reads what the other writes)
    cudaStream_t st_sort = nullptr;
    cudaEvent_t ev_push = nullptr, ev_sort = nullptr;
    bool sort_pending = false;     // pa[cur][0..n) is in arbitrary order (left by the pipelined host-buffer step): sort before use
    bool defer_count = false;      // slabs, peer-ring step: count_moves is left to the sort stream (mpic_peer_step_end)
    bool moves_counted = false;    // cell_count already holds the destination histogram of the last push (count_moves ran)
    cudaEvent_t pipe_ev_up[MPIC_PIPE_MAX] = {}, pipe_ev_done[MPIC_PIPE_MAX] = {}, pipe_ev_free = nullptr;
    int* pipe_count = nullptr;     // per-chunk cell histogram / scatter cursor
    int* pipe_start = nullptr;     // per-chunk cell offsets (relative to the chunk's first slot)
    std::vector<int> owner;        // global tiles
    // device scratch of the tile-layout converters and the occupancy queries (grown on demand, reused: no cudaMalloc per call)
    void* tile_tmp = nullptr;   size_t tile_tmp_bytes = 0;     // tile lists / counts / per-particle tile ids
    void* tile_stage = nullptr; size_t tile_stage_bytes = 0;   // tile patches with guards
    // run-time re-partition (reslab.cu): parcels this handle packed and exported, freed by mpic_reslab_finish
    std::vector<void*> parcels;
    int64_t kept_cap = 0;          // capacity of the particle arrays kept across mpic_free_state (0: none)
    // asynchronous field dumps (dump.cu): side stream, pinned staging, writer thread
    struct DumpState* dump = nullptr;
    void* ring_events = nullptr;   // ring.cu: per-handle events of the single-process peer-copy ring
};

void mpic_ring_release(mpic_sim* s);
int mpic_set_geometry(mpic_sim* s);                  // api.cu: G, P, cf from prm
int mpic_alloc_state(mpic_sim* s);                   // api.cu: every device array sized by the slab rows / particle capacity
void mpic_free_state(mpic_sim* s, int64_t keep_cap); // api.cu: ... and back (releases the peer / ring mappings first); keep_cap > 0:
                                                     // keep the capacity-sized particle arrays if the capacity is exactly this
int mpic_scratch(void** buf, size_t* have, size_t need);   // api.cu: grow-only device scratch
extern "C" int mpic_ensure_sorted(mpic_sim* s);     // api.cu: run the deferred sort of the pipelined host-buffer step, if any
extern "C" int mpic_sort_uploaded(mpic_sim* s);      // api.cu: sort_all(false) -- keys from the stored positions
int mpic_refresh_x_guards(mpic_sim* s);              // api.cu: periodic x images (and y on a single slab) of the six planes
int mpic_fused_setup(mpic_sim* s);                   // api.cu: second plane block + tensor maps (idempotent)
void mpic_fused_swap(mpic_sim* s);                   // api.cu: planes <-> planes_alt (+ maps, + peer pointers)
void mpic_peer_release(mpic_sim* s);
extern "C" int mpic_sort_after_push(mpic_sim* s);   // api.cu: sort_all(counted)
extern "C" int mpic_sort_after_push_on(mpic_sim* s, void* stream);   // the same, every launch on `stream`; the main stream then waits for it
#define HALO_HDR 16                /* values (128 B) of header in front of each halo message; [0] = leaver count to that side */
#define REQUIRE(cond, code, ...) do { if (!(cond)) { mpic_set_error(__VA_ARGS__); return code; } } while (0)
#define TRY(expr) do { int _r = (expr); if (_r) return _r; } while (0)

