// common.cuh -- shared device/host definitions for libminipic_b200.so (sm_100a only).
//
// HBM layout (DESIGN.md "data layout"):
//   fields   six FP64 planes Ex,Ey,Ez,Bx,By,Bz + three int64 fixed-point current accumulators Jx,Jy,Jz, each
//            (nyl + 2g) rows x pitch values.  Interior cell (i,j) of the slab lives at row g+j, column X0+i with
//            X0 = 16 values (128 B) so interior rows start on a 128-byte line; pitch is a multiple of 16 values.
//            The guard ring (g=2 on every side) always holds the periodic image (x: local wrap; y: local wrap on one
//            GPU, neighbour slab rows otherwise), i.e. exactly what the reference's tiles hold after an exchange phase
//            (PIC_2D.cpp:200-248) -- so any tile-with-guards patch is a plain 2-D window of the planes.
//   particles SoA x,y,ux,uy,uz,q (FP64) + tag (int32), kept sorted by key = local_tile*cells_per_tile + cell_in_tile;
//            cell_start[key] (int32) gives each cell's first slot, so a tile's particles are one contiguous range.
#pragma once
#include <cuda_runtime.h>
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/minipic_b200.h"

#define MPIC_X0 16
#define MPIC_MAX_DEVICES 64                             /* per-device caches of function attributes */
#define MPIC_JSCALE 4503599627370496.0               /* 2^52: fixed-point scale of the current accumulators */
#define MPIC_INV_JSCALE (1.0 / 4503599627370496.0)

struct FieldGeom {
    int nx, nyl;          // interior cells of this slab
    int g;                // guard
    int pitch;            // values per row
    int rows;             // nyl + 2g
    int64_t plane;        // values per plane = rows*pitch
    int ywrap;            // 1: y guards are the local periodic image (single slab)
    unsigned int* err;    // device error word of the handle (MPIC_ERR_* bits), read back by mpic_sync
};
#define MPIC_ERR_POSITION 1u   /* a stored particle position was outside [0,box) or not finite (upload / arrival) */
#define MPIC_ERR_JRANGE 2u     /* a fixed-point current accumulator came within a factor 2 of the int64 range (|J| >= 1024) */

// Peer-memory halo: where the Yee sweeps store the images of their edge rows in the two ring neighbours' guard rows
// (device pointers into the neighbours' 6-plane blocks, mapped by peer.cu).  All null: single slab or staged transport.
struct PeerHalo {
    double* lo = nullptr;     // lower neighbour's planes; my interior rows [0,g) -> its rows g+nyl_lo+j (its upper guard)
    double* hi = nullptr;     // upper neighbour's planes; my interior rows [nyl-g,nyl) -> its rows j-(nyl-g) (its lower guard)
    int64_t lo_plane = 0, hi_plane = 0;
    int lo_row0 = 0;          // g + nyl_lo
    // fused PIC step on a slab: the current of the cell row just above the slab (needed for the ring of E') is the upper
    // neighbour's first interior row of ITS accumulators (complete once both ranks' hand-offs are done), read over NVLink
    const int64_t* hi_j = nullptr;
    int64_t hi_jplane = 0;
};

// Phase ordering fused into the sweeps of a peer-memory ring (peer.cu): only the CTAs that own slab-edge rows wait for the
// neighbour's flag (before touching its guard rows or reading their own), the interior CTAs start at once -- the exchange
// latency hides behind the interior of the sweep; the last edge CTA of each side raises this sweep's flag at the neighbour.
struct PeerSync {
    const unsigned long long* wait_flags = nullptr;   // own control words of the awaited phase: [0] lower, [1] upper neighbour
    unsigned long long want = 0;                      // 0: nothing to await
    const unsigned long long* wait_local = nullptr;   // an additional LOCAL word every edge CTA awaits (own hand-off finished)
    unsigned long long want_local = 0;
    unsigned long long* sig_lo = nullptr;             // this sweep's flag word at the lower / upper neighbour
    unsigned long long* sig_hi = nullptr;
    unsigned long long sig = 0;
    unsigned int* done = nullptr;                     // [2] edge-CTA counters (self-resetting); null: no fused ordering
    unsigned int n_lo = 0, n_hi = 0;                  // edge CTAs per side
    unsigned long long* err = nullptr;
    unsigned long long timeout_ns = 0;
    int phase = 0;
};

struct PicConst {
    int nx, ny;           // GLOBAL cells
    int tnx, tny;         // tile interior
    int gtc;              // global tile columns
    int ltr;              // local tile rows
    int cy0;              // first global cell row of this slab (row0*tny)
    int nyl;              // local cell rows
    int cpt;              // cells per tile
    int ncells;           // local cells
    double dx, dy, dt;
    double inv_dx, inv_dy;
    double box_x, box_y;
    double h;             // 0.5*dt*qm
    double cjx, cjy, cjz; // 1/(dy*dt), 1/(dx*dt), 1/(dx*dy)
    double sjx, sjy, sjz; // the same times 2^52 (exact): fixed-point scale folded into the charge factor
    double dnx, dny;      // (double)nx, (double)ny
    int deposit;          // 0: charge-conserving zigzag (default); 1: direct CIC (SURVEY.md B.5 option a)
};

struct Particles {
    double *x, *y, *ux, *uy, *uz, *q;
    int* tag;
};

#define MPIC_REC_DOUBLES 7                            /* migration record: x,y,ux,uy,uz,q,tag(as int64 bits) = 56 B */
struct LeaveBufs {
    double* rec[2];       // AoS records of particles that left the slab downwards (0) / upwards (1)
    int* count;           // count[2]
    int cap;              // records per buffer
    // peer-memory ring: the CTAs of the slab's first / last tile row await the neighbours' last field phase (their stores
    // into this slab's guard rows) before staging their field patch -- no separate wait kernel in front of the push
    const unsigned long long* wait_flags = nullptr;   // [0] lower, [1] upper neighbour
    unsigned long long wait_want = 0;                 // 0: nothing to await
    unsigned long long* wait_err = nullptr;
    unsigned long long wait_timeout_ns = 0;
    int wait_phase = 0;
};

void mpic_set_error(const char* fmt, ...);
// Identity of this process for "may I use the raw device pointers in this blob?": the pid alone is not enough (ranks in separate
// PID namespaces -- one container per rank -- can share a pid), so blobs also carry a random 64-bit value drawn once per library
// load, and both must match.
unsigned long long mpic_process_nonce();

#define CUDA_TRY(expr)                                                                          \
    do {                                                                                        \
        cudaError_t _e = (expr);                                                                \
        if (_e != cudaSuccess) {                                                                \
            cudaGetLastError(); /* reported here: it must not surface again at a later launch check */ \
            mpic_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return _e == cudaErrorNoDevice || _e == cudaErrorInsufficientDriver ? MPIC_ENODEV : MPIC_ECUDA; \
        }                                                                                       \
    } while (0)

static inline int ceil_div(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// ---- kernel launchers implemented in yee.cu
struct YeeCoef { double bx, by, ex, ey, dt; int fma; };   // bx = dt/(2dx) or dt/dx ; ex = dt/dx
int launch_b_sweep(const FieldGeom& G, double* planes, const YeeCoef& c, const PeerHalo& ph, PeerSync ps, cudaStream_t st);
int launch_e_sweep(const FieldGeom& G, double* planes, int64_t* jacc, int with_j, const YeeCoef& c, const PeerHalo& ph, PeerSync ps, cudaStream_t st);
int launch_fused_step(const FieldGeom& G, const CUtensorMap* src_map, double* dst, const int64_t* jacc, int with_j, const YeeCoef& c,
                      const PeerHalo& dst_halo, PeerSync ps, cudaStream_t st);
void fused_box_dims(int* pw, int* ph);
int launch_wrap_guards(const FieldGeom& G, double* planes, int first_plane, int nplanes, int do_y, cudaStream_t st);
int launch_fold_current(const FieldGeom& G, int64_t* jacc, int nplanes, int do_y, cudaStream_t st);
// slab halos: staging layout [side][3 planes][g rows][pitch] (+ header), side 0 = lower neighbour, 1 = upper neighbour
int launch_halo_pack(const FieldGeom& G, const double* base3, int is_current, double* send, int64_t side_stride, cudaStream_t st);
int launch_halo_unpack(const FieldGeom& G, double* base3, int is_current, const double* recv, int64_t side_stride, cudaStream_t st);
int launch_planes_to_tiles(const FieldGeom& G, const double* planes, double* grid6, const int* lrow, const int* lcol,
                           int ntiles, int tnx, int tny, cudaStream_t st);
int launch_tiles_to_planes(const FieldGeom& G, double* planes, const double* grid6, const int* lrow, const int* lcol,
                           int ntiles, int tnx, int tny, cudaStream_t st);
int launch_tile_b_half(const double* E3, const double* Bo3, double* Bn3, int tnx, int tny, int g, double cx, double cy, cudaStream_t st);
int launch_tile_e_full(double* E3, const double* B3, int tnx, int tny, int g, double cx, double cy, cudaStream_t st);
int launch_strip_copy(const double* src6, double* dst6, int TX, int r0, int c0, int nr, int nc, int to_strip, cudaStream_t st);
int launch_field_energy(const FieldGeom& G, const double* planes, double* out, double* work, cudaStream_t st);   // work: MPIC_DIAG_WORK doubles
int launch_gauss_residual(const FieldGeom& G, const double* planes, const int64_t* rho, double mean_rho, double inv_dx,
                          double inv_dy, double* out_max, int j_first, cudaStream_t st);

// ---- kernel launchers implemented in particles.cu
int launch_push_deposit(const PicConst& P, const FieldGeom& G, const CUtensorMap* tmap, const CUtensorMap* tmap_pack, int64_t* jacc,
                        Particles src, Particles dst, const int* perm, const int* cell_start, int* mcnt, unsigned char* dirs, int split,
                        LeaveBufs lv, cudaStream_t st, int64_t n_hint);   // n_hint: particles in the slab (picks the sparse-cell variant)
int launch_count_moves(const PicConst& P, int ywrap, const int* mcnt, int* cell_count, cudaStream_t st);
int launch_offsets(const PicConst& P, int ywrap, const int* mcnt, const int* start_new, int* offtab, int* cursor, cudaStream_t st);
int launch_perm_stable(const PicConst& P, int ywrap, const int* cs_old, const unsigned char* dirs, const int* offtab, int* perm,
                       int64_t n_hint, cudaStream_t st);
int launch_arrivals(const PicConst& P, const int* keys, int64_t first, int n, int* cursor, const int* start_new, int* perm, cudaStream_t st);
int launch_leavers(const PicConst& P, Particles o, const int* cs_old, const unsigned char* dirs, const int* mcnt, int* lv_off,
                   int* cell_count, LeaveBufs lv, cudaStream_t st);
int launch_segsort(const int* start, int ncells, int* perm, int* scratch, cudaStream_t st);
int launch_perm_build(const PicConst& P, const int* keys, int64_t n, int* cursor, int* perm, cudaStream_t st);
int launch_gather(Particles src, Particles dst, int* perm, int64_t n, const int* n_dev, cudaStream_t st);
int launch_absorb(const PicConst& P, Particles p, int64_t first, const double* rec, int n, int* cell_count, int* keys,
                  unsigned int* err, cudaStream_t st);
int launch_count_keys(const PicConst& P, Particles p, int64_t first, int64_t n, int* cell_count, int* keys, unsigned int* err, cudaStream_t st);
int launch_scan_cells(int* cell_count, int* cell_start, int n_entries, int* block_tmp, cudaStream_t st);
int launch_crossers(const unsigned char* dirs, int64_t n, unsigned long long* out, cudaStream_t st);
int launch_tile_counts(const PicConst& P, const int* cell_start, int* counts, cudaStream_t st);
int launch_particle_tiles(const PicConst& P, Particles p, int64_t n, int row0, int* gid, cudaStream_t st);
int launch_kinetic(Particles p, int64_t n, double* out2, double* work, cudaStream_t st);
#define MPIC_DIAG_WORK (2 + 296 + 2 * 592)             /* two tickets + block partials of the energy reductions */
int launch_deposit_rho(const PicConst& P, const FieldGeom& G, Particles p, int64_t n, int64_t* rho, cudaStream_t st);
int launch_load_uniform(const PicConst& P, Particles p, int64_t first, int64_t nsel, int cy_lo, int64_t tag0, int ppc,
                        double u_th, double d0, double d1, double d2, double two_stream, double qp, uint64_t seed,
                        cudaStream_t st);
size_t push_smem_bytes(int tnx, int tny);
void push_box_dims(int tnx, int tny, int* pw, int* ph);
void pack_box_dims(int tnx, int tny, int* pw, int* ph);   // packed-walk kernel: (PW, PACK_ROWS + 2)
int launch_sum_i64(const int64_t* a, int64_t n, double* out, cudaStream_t st);
int launch_selftest_math(int64_t n, uint64_t seed, unsigned long long* bad, cudaStream_t st);
int yee_preload();
int particles_preload();
