// api.cu -- the C ABI (include/minipic_b200.h): handle life cycle, host<->device staging, phase sequencing.
// Phase order of one PIC step (extending PIC_2D.cpp:171-420):
//   P1/P2 push+deposit -> P3 J guard fold -> A half-B (+guards) -> C E sweep with -dt*J (+guards) -> E half-B (+guards)
//   -> S per-cell count (fused in the push) / scan / scatter.
#include "common.cuh"
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>
#include <math.h>
#include <vector>
#include <string>
#include <utility>
#include <thread>

static thread_local char g_err[512] = "";
void mpic_set_error(const char* fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof g_err, fmt, ap); va_end(ap);
}

#include <unistd.h>
#include <time.h>
#include <mutex>
unsigned long long mpic_process_nonce() {
    static std::once_flag once;
    static unsigned long long nonce = 0;
    std::call_once(once, [] {
        unsigned long long v = 0;
        if (FILE* f = fopen("/dev/urandom", "rb")) { if (fread(&v, sizeof v, 1, f) != 1) v = 0; fclose(f); }
        if (!v) {                                         // no urandom: clock, pid and an address of this mapping, mixed
            struct timespec ts; clock_gettime(CLOCK_REALTIME, &ts);
            v = (unsigned long long)ts.tv_nsec * 0x9E3779B97F4A7C15ull ^ (unsigned long long)ts.tv_sec << 20 ^
                (unsigned long long)getpid() << 44 ^ (unsigned long long)(uintptr_t)&nonce;
        }
        nonce = v | 1ull;                                 // never zero: a zeroed blob matches nobody
    });
    return nonce;
}

#include "sim.cuh"
static void dump_release(mpic_sim* s);

extern "C" {

const char* mpic_last_error(void) { return g_err; }
int mpic_abi_version(void) { return 3; }
int mpic_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// ------------------------------------------------------------------------------------------ integer helpers
static const int k_drow[8] = { 0, 0, -1, 1, -1, -1, 1, 1 };
static const int k_dcol[8] = { -1, 1, 0, 0, -1, 1, -1, 1 };
static const int k_opp[8] = { 1, 0, 3, 2, 7, 6, 5, 4 };
static void best_grid(int n, int* a, int* b) {
    int r = (int)sqrt((double)n);
    while (r > 1 && n % r != 0) r--;
    if (r < 1) r = 1;
    *a = r; *b = n / r;
}
void mpic_find_best_grid(int size, int* R, int* C) { best_grid(size, R, C); }
void mpic_find_best_tile_grid(int n, int* tr, int* tc) { best_grid(n, tr, tc); }
int mpic_tile_global_row(int rank_row, int tile_rows, int t_row) { return rank_row * tile_rows + t_row; }
int mpic_tile_global_col(int rank_col, int tile_cols, int t_col) { return rank_col * tile_cols + t_col; }
int mpic_get_global_id(int row, int col, int gtc) { return row * gtc + col; }
void mpic_get_global_row_col(int gid, int gtc, int* row, int* col) { *row = gid / gtc; *col = gid % gtc; }
int mpic_get_neighbor_gid(int gid, int d, int gtr, int gtc) {
    int row = gid / gtc + k_drow[d & 7], col = gid % gtc + k_dcol[d & 7];
    row = row < 0 ? gtr - 1 : row >= gtr ? 0 : row;
    col = col < 0 ? gtc - 1 : col >= gtc ? 0 : col;
    return row * gtc + col;
}
int mpic_compute_tag(int gid, int d) { return gid * 8 + d; }
void mpic_direction_tables(int drow[8], int dcol[8], int opposite[8]) {
    for (int d = 0; d < 8; d++) { drow[d] = k_drow[d]; dcol[d] = k_dcol[d]; opposite[d] = k_opp[d]; }
}
void mpic_slab_rows(int gtr, int nranks, int rank, int* row0, int* nrows) {
    int base = gtr / nranks, rem = gtr % nranks;
    *row0 = rank * base + (rank < rem ? rank : rem);
    *nrows = base + (rank < rem ? 1 : 0);
}
int mpic_strip_cells(int d, int tnx, int tny, int g) { return d < 2 ? tny * g : d < 4 ? tnx * g : g * g; }

static void strip_geom(int d, int tnx, int tny, int g, bool interior, int* r0, int* c0, int* nr, int* nc) {
    const int cs = k_dcol[d], rs = k_drow[d];
    if (cs == 0) { *c0 = g; *nc = tnx; } else { *nc = g; *c0 = interior ? (cs < 0 ? g : tnx) : (cs < 0 ? 0 : tnx + g); }
    if (rs == 0) { *r0 = g; *nr = tny; } else { *nr = g; *r0 = interior ? (rs < 0 ? g : tny) : (rs < 0 ? 0 : tny + g); }
}

// ------------------------------------------------------------------------------------------ L0 compat
static int need_device() {
    REQUIRE(mpic_device_count() > 0, MPIC_ENODEV, "no CUDA device: libminipic_b200 has no CPU fallback");
    return 0;
}
int mpic_update_b_half(const double* E3, const double* Bold3, double* Bnew3, int tnx, int tny, int g, double dt, double dx, double dy) {
    REQUIRE(E3 && Bold3 && Bnew3 && tnx > 0 && tny > 0 && g >= 1, MPIC_EINVAL, "mpic_update_b_half: bad argument");
    TRY(need_device());
    const size_t bytes = (size_t)(tnx + 2 * g) * (tny + 2 * g) * 24;
    double *dE, *dBo, *dBn;
    CUDA_TRY(cudaMalloc(&dE, 3 * bytes));
    dBo = dE + bytes / 8; dBn = dBo + bytes / 8;
    CUDA_TRY(cudaMemcpy(dE, E3, bytes, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(dBo, Bold3, bytes, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(dBn, Bnew3, bytes, cudaMemcpyHostToDevice));   // guards of BhalfNew stay as the caller set them
    int r = launch_tile_b_half(dE, dBo, dBn, tnx, tny, g, dt / (2 * dx), dt / (2 * dy), 0);
    if (!r) { cudaError_t e = cudaMemcpy(Bnew3, dBn, bytes, cudaMemcpyDeviceToHost); if (e != cudaSuccess) { mpic_set_error("D2H: %s", cudaGetErrorString(e)); r = MPIC_ECUDA; } }
    cudaFree(dE);
    return r;
}
int mpic_update_e_full(double* E3, const double* B3, int tnx, int tny, int g, double dt, double dx, double dy) {
    REQUIRE(E3 && B3 && tnx > 0 && tny > 0 && g >= 1, MPIC_EINVAL, "mpic_update_e_full: bad argument");
    TRY(need_device());
    const size_t bytes = (size_t)(tnx + 2 * g) * (tny + 2 * g) * 24;
    double *dE, *dB;
    CUDA_TRY(cudaMalloc(&dE, 2 * bytes));
    dB = dE + bytes / 8;
    CUDA_TRY(cudaMemcpy(dE, E3, bytes, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(dB, B3, bytes, cudaMemcpyHostToDevice));
    int r = launch_tile_e_full(dE, dB, tnx, tny, g, dt / dx, dt / dy, 0);
    if (!r) { cudaError_t e = cudaMemcpy(E3, dE, bytes, cudaMemcpyDeviceToHost); if (e != cudaSuccess) { mpic_set_error("D2H: %s", cudaGetErrorString(e)); r = MPIC_ECUDA; } }
    cudaFree(dE);
    return r;
}
static int strip_op(double* grid6, int tnx, int tny, int g, int d, double* strip6, bool pack) {
    REQUIRE(grid6 && strip6 && d >= 0 && d < 8 && tnx >= g && tny >= g && g >= 1, MPIC_EINVAL, "strip: bad argument");
    TRY(need_device());
    int r0, c0, nr, nc;
    strip_geom(d, tnx, tny, g, pack, &r0, &c0, &nr, &nc);
    const int TX = tnx + 2 * g, TY = tny + 2 * g;
    const size_t gb = (size_t)TX * TY * 48, sb = (size_t)nr * nc * 48;
    double *dg, *ds;
    CUDA_TRY(cudaMalloc(&dg, gb + sb));
    ds = dg + gb / 8;
    CUDA_TRY(cudaMemcpy(dg, grid6, gb, cudaMemcpyHostToDevice));
    if (!pack) CUDA_TRY(cudaMemcpy(ds, strip6, sb, cudaMemcpyHostToDevice));
    int r = pack ? launch_strip_copy(dg, ds, TX, r0, c0, nr, nc, 1, 0) : launch_strip_copy(ds, dg, TX, r0, c0, nr, nc, 0, 0);
    if (!r) {
        cudaError_t e = pack ? cudaMemcpy(strip6, ds, sb, cudaMemcpyDeviceToHost) : cudaMemcpy(grid6, dg, gb, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { mpic_set_error("D2H: %s", cudaGetErrorString(e)); r = MPIC_ECUDA; }
    }
    cudaFree(dg);
    return r ? r : nr * nc;
}
int mpic_pack_send_buffer(const double* grid6, int tnx, int tny, int g, int d, double* out6) {
    return strip_op(const_cast<double*>(grid6), tnx, tny, g, d, out6, true);
}
int mpic_update_guard_region(double* grid6, int tnx, int tny, int g, int d, const double* in6) {
    int r = strip_op(grid6, tnx, tny, g, d, const_cast<double*>(in6), false);
    return r < 0 ? r : 0;
}

// ------------------------------------------------------------------------------------------ handle
static int encode_box_map(mpic_sim* s, double* planes, int pw, int ph, CUtensorMap* out) {
    typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CUDA_TRY(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    REQUIRE(fn && qres == cudaDriverEntryPointSuccess, MPIC_ECUDA, "cuTensorMapEncodeTiled not available");
    REQUIRE(pw <= 256 && ph <= 256, MPIC_EINVAL, "tile too large for one TMA box (max 254 cells per side)");
    cuuint64_t gdim[3] = { (cuuint64_t)s->G.pitch, (cuuint64_t)s->G.rows, 6 };
    cuuint64_t gstr[2] = { (cuuint64_t)s->G.pitch * 8, (cuuint64_t)s->G.plane * 8 };
    cuuint32_t box[3] = { (cuuint32_t)pw, (cuuint32_t)ph, 6 };
    cuuint32_t estr[3] = { 1, 1, 1 };
    CUresult r = ((encode_fn)fn)(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, planes, gdim, gstr, box, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    REQUIRE(r == CUDA_SUCCESS, MPIC_ECUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
    return 0;
}
static int make_tensor_map(mpic_sim* s) {
    int pw, ph;
    push_box_dims(s->P.tnx, s->P.tny, &pw, &ph);
    TRY(encode_box_map(s, s->planes, pw, ph, &s->tmap));
    pack_box_dims(s->P.tnx, s->P.tny, &pw, &ph);
    TRY(encode_box_map(s, s->planes, pw, ph, &s->tmap_pack));
    s->has_tmap = true;
    return 0;
}

static int alloc_particles(Particles* p, int64_t cap) {
    double* base;
    CUDA_TRY(cudaMalloc(&base, (size_t)cap * 6 * 8));
    p->x = base; p->y = base + cap; p->ux = base + 2 * cap; p->uy = base + 3 * cap; p->uz = base + 4 * cap; p->q = base + 5 * cap;
    CUDA_TRY(cudaMalloc(&p->tag, (size_t)cap * 4));
    return 0;
}

// ---- handle state in three layers: (1) geometry = everything derived from prm (set_geometry), (2) device state sized by the
// slab's rows and the particle capacity (alloc_state / free_state), (3) streams, events, owner map, counters (mpic_create /
// mpic_destroy).  mpic_reslab_apply (reslab.cu) re-partitions a LIVE handle by replacing layers 1 and 2 only.
}  // extern "C"
static int validate_params(const mpic_params* p) {
    REQUIRE(p->guard == 2, MPIC_EINVAL, "guard must be 2 (PIC_2D.cpp:65)");
    REQUIRE(p->nx > 0 && p->ny > 0 && p->tnx > 0 && p->tny > 0 && p->nx % p->tnx == 0 && p->ny % p->tny == 0, MPIC_EINVAL,
            "nx, ny must be positive multiples of the tile size (the reference silently drops cells otherwise, PIC_2D.cpp:67-68)");
    REQUIRE(p->tnx >= 2 * p->guard && p->tny >= 2 * p->guard, MPIC_EINVAL, "tile interior must be >= 2*guard cells");
    REQUIRE(p->dx > 0 && p->dy > 0 && p->dt > 0, MPIC_EINVAL, "dx, dy, dt must be positive");
    // the push (and its shared-memory count / current patches) assumes a particle moves less than one cell per step: |v| < c = 1
    REQUIRE(p->max_particles == 0 || (p->dt <= p->dx && p->dt <= p->dy), MPIC_EINVAL,
            "dt = %g exceeds min(dx, dy) = %g: a particle could cross more than one cell per step", p->dt, p->dx < p->dy ? p->dx : p->dy);
    REQUIRE(p->nranks >= 1 && p->rank >= 0 && p->rank < p->nranks, MPIC_EINVAL, "bad rank/nranks");
    REQUIRE(p->deposit == 0 || p->deposit == 1, MPIC_EINVAL, "deposit must be 0 (zigzag) or 1 (direct CIC)");
    const int gtr = p->ny / p->tny;
    REQUIRE(p->row0 >= 0 && p->nrows >= 1 && p->row0 + p->nrows <= gtr, MPIC_EINVAL, "bad slab rows");
    REQUIRE(p->nranks > 1 || (p->row0 == 0 && p->nrows == gtr), MPIC_EINVAL, "single rank must own every tile row");
    // 32-bit slot / cell indices and int32 tags (mpic_load_uniform numbers particles globally: 1.07 B fits, 2^31 does not)
    REQUIRE((int64_t)p->nx * p->nrows * p->tny < (1ll << 31) - 8 && p->max_particles >= 0 && p->max_particles < (1ll << 31) - 64, MPIC_EINVAL,
            "slab too large for 32-bit indices (cells and max_particles must stay below 2^31)");
    return 0;
}
int mpic_set_geometry(mpic_sim* s) {
    const mpic_params* p = &s->prm;
    const int gtc = p->nx / p->tnx;
    FieldGeom& G = s->G;
    G.nx = p->nx; G.nyl = p->nrows * p->tny; G.g = p->guard;
    G.pitch = (MPIC_X0 + p->nx + p->guard + 15) / 16 * 16;
    G.rows = G.nyl + 2 * p->guard;
    G.plane = (int64_t)G.rows * G.pitch;
    G.ywrap = p->nranks == 1;
    PicConst& P = s->P;
    P.nx = p->nx; P.ny = p->ny; P.tnx = p->tnx; P.tny = p->tny; P.gtc = gtc; P.ltr = p->nrows; P.cy0 = p->row0 * p->tny;
    P.nyl = G.nyl; P.cpt = p->tnx * p->tny; P.ncells = p->nx * G.nyl;
    P.dx = p->dx; P.dy = p->dy; P.dt = p->dt; P.inv_dx = 1.0 / p->dx; P.inv_dy = 1.0 / p->dy;
    P.box_x = p->nx * p->dx; P.box_y = p->ny * p->dy;
    P.h = 0.5 * p->dt * p->qm;
    P.cjx = 1.0 / (p->dy * p->dt); P.cjy = 1.0 / (p->dx * p->dt); P.cjz = 1.0 / (p->dx * p->dy);
    P.sjx = P.cjx * MPIC_JSCALE; P.sjy = P.cjy * MPIC_JSCALE; P.sjz = P.cjz * MPIC_JSCALE; P.dnx = (double)p->nx; P.dny = (double)p->ny;
    P.deposit = p->deposit;
    YeeCoef& c = s->cf;
    c.bx = p->scheme ? p->dt / p->dx : p->dt / (2 * p->dx);
    c.by = p->scheme ? p->dt / p->dy : p->dt / (2 * p->dy);
    c.ex = p->dt / p->dx; c.ey = p->dt / p->dy; c.dt = p->dt; c.fma = p->fma_mode;
    return 0;
}
int mpic_alloc_state(mpic_sim* s) {
    const mpic_params* p = &s->prm;
    FieldGeom& G = s->G;
    const PicConst& P = s->P;
    CUDA_TRY(cudaMalloc(&s->planes, (size_t)6 * G.plane * 8));
    CUDA_TRY(cudaMemsetAsync(s->planes, 0, (size_t)6 * G.plane * 8, s->st));
    CUDA_TRY(cudaMalloc(&s->jacc, (size_t)3 * G.plane * 8));
    CUDA_TRY(cudaMemsetAsync(s->jacc, 0, (size_t)3 * G.plane * 8, s->st));
    CUDA_TRY(cudaMalloc(&s->d_scratch, (size_t)(16 + MPIC_DIAG_WORK) * 8));   // [0,8) results, [8] error word, [16..) reduction work space
    G.err = (unsigned int*)(s->d_scratch + 8);            // the handle's device error word
    CUDA_TRY(cudaMemsetAsync(s->d_scratch, 0, (size_t)(16 + MPIC_DIAG_WORK) * 8, s->st));
    s->cap = p->max_particles;
    if (s->cap > 0) {
        if (!(s->kept_cap == s->cap && s->pa[0].x)) {        // (a re-partition keeps the capacity-sized arrays when the capacity stays)
            TRY(alloc_particles(&s->pa[0], s->cap));
            TRY(alloc_particles(&s->pa[1], s->cap));
            CUDA_TRY(cudaMalloc(&s->keys, (size_t)s->cap * 4));
            CUDA_TRY(cudaMalloc(&s->perm, (size_t)s->cap * 4));
            CUDA_TRY(cudaMalloc(&s->dirs, (size_t)s->cap));
        }
        s->kept_cap = 0;
        s->n_entries = P.ncells + 2;
        CUDA_TRY(cudaMalloc(&s->cell_start, (size_t)(s->n_entries + 1) * 4));
        CUDA_TRY(cudaMalloc(&s->cell_count, (size_t)s->n_entries * 4));
        CUDA_TRY(cudaMalloc(&s->scan_tmp, (size_t)(s->n_entries / 2048 + 2) * 4));
        CUDA_TRY(cudaMalloc(&s->cell_start_alt, (size_t)(s->n_entries + 1) * 4));
        CUDA_TRY(cudaMalloc(&s->mcnt, (size_t)9 * P.ncells * 4));
        CUDA_TRY(cudaMalloc(&s->offtab, (size_t)9 * P.ncells * 4));
        CUDA_TRY(cudaMemsetAsync(s->mcnt, 0, (size_t)9 * P.ncells * 4, s->st));
        CUDA_TRY(cudaMemsetAsync(s->cell_start, 0, (size_t)(s->n_entries + 1) * 4, s->st));
        CUDA_TRY(cudaMemsetAsync(s->cell_start_alt, 0, (size_t)(s->n_entries + 1) * 4, s->st));
        if (p->nranks > 1) CUDA_TRY(cudaMalloc(&s->lv_off, (size_t)2 * p->nx * 4));
        TRY(make_tensor_map(s));
        // enough CTAs per tile that the grid covers >= 4 waves of 2 CTAs/SM even for few tiles
        int ntiles = P.ltr * P.gtc;
        s->push_split = 1;
        while (ntiles * s->push_split < 12 * s->nsm && s->push_split < 64) s->push_split *= 2;   // >= 4 waves of 3 CTAs/SM
        if (const char* e = getenv("MPIC_PUSH_SPLIT")) { int v = atoi(e); if (v >= 1 && v <= 256) s->push_split = v; }   // tuning knob
    }
    if (p->nranks > 1) {
        s->halo_side = HALO_HDR + (int64_t)3 * G.g * G.pitch;
        CUDA_TRY(cudaMalloc(&s->halo_send, (size_t)2 * s->halo_side * 8));
        CUDA_TRY(cudaMalloc(&s->halo_recv, (size_t)2 * s->halo_side * 8));
        CUDA_TRY(cudaMemsetAsync(s->halo_send, 0, (size_t)2 * s->halo_side * 8, s->st));
        CUDA_TRY(cudaMemsetAsync(s->halo_recv, 0, (size_t)2 * s->halo_side * 8, s->st));
        if (s->cap > 0) {
            int64_t lc = 4 * (s->cap / G.nyl + 1);
            if (lc < 4096) lc = 4096;
            if (lc > s->cap) lc = s->cap;
            s->lv.cap = (int)lc;
            // arrival block: [2 step parities][2 sides][lc records]; the staged transports use parity 0
            CUDA_TRY(cudaMalloc(&s->arrive_blk, (size_t)4 * lc * MPIC_REC_DOUBLES * 8));
            for (int d = 0; d < 2; d++) {
                CUDA_TRY(cudaMalloc(&s->lv.rec[d], (size_t)lc * MPIC_REC_DOUBLES * 8));
                s->arrive[d] = s->arrive_blk + (size_t)d * lc * MPIC_REC_DOUBLES;
            }
        }
    }
    CUDA_TRY(cudaMalloc(&s->lv.count, 2 * sizeof(int)));
    CUDA_TRY(cudaMemsetAsync(s->lv.count, 0, 2 * sizeof(int), s->st));
    return 0;
}
// frees layer 2 and resets every flag that describes it (the peer / ring mappings into OTHER handles are released first; the
// caller guarantees that no other handle still maps THIS one's memory)
void mpic_free_state(mpic_sim* s, int64_t keep_cap) {
    mpic_ring_release(s);
    mpic_peer_release(s);
    // capacity-sized arrays (both particle buffers, keys, perm, dirs: nearly all of the memory) survive a re-partition that
    // keeps the capacity: freeing and re-allocating tens of GB dominated the first measured rebalance (profiles/r02af)
    const bool keep = keep_cap > 0 && keep_cap == s->cap && s->pa[0].x;
    Particles kp[2] = { s->pa[0], s->pa[1] };
    int *kkeys = s->keys, *kperm = s->perm;
    unsigned char* kdirs = s->dirs;
    if (keep) { s->pa[0] = s->pa[1] = Particles(); s->keys = s->perm = nullptr; s->dirs = nullptr; }
    cudaFree(s->planes); cudaFree(s->planes_alt); cudaFree(s->jacc); cudaFree(s->rho); cudaFree(s->d_scratch);
    s->planes = s->planes_alt = nullptr; s->jacc = s->rho = nullptr; s->d_scratch = nullptr;
    for (int k = 0; k < 2; k++) { cudaFree(s->pa[k].x); cudaFree(s->pa[k].tag); s->pa[k] = Particles(); }
    cudaFree(s->cell_start); cudaFree(s->cell_count); cudaFree(s->scan_tmp); cudaFree(s->keys); cudaFree(s->perm);
    cudaFree(s->cell_start_alt); cudaFree(s->mcnt); cudaFree(s->offtab); cudaFree(s->dirs); cudaFree(s->lv_off);
    s->cell_start = s->cell_count = s->scan_tmp = s->keys = s->perm = s->cell_start_alt = s->mcnt = s->offtab = s->lv_off = nullptr;
    s->dirs = nullptr;
    cudaFree(s->halo_send); cudaFree(s->halo_recv); cudaFree(s->lv.count);
    s->halo_send = s->halo_recv = nullptr;
    for (int d = 0; d < 2; d++) { cudaFree(s->lv.rec[d]); s->arrive[d] = nullptr; }
    cudaFree(s->arrive_blk);
    s->arrive_blk = nullptr;
    s->lv = LeaveBufs();
    cudaFree(s->pipe_count); cudaFree(s->pipe_start);
    s->pipe_count = s->pipe_start = nullptr;
    cudaFree(s->tile_tmp); cudaFree(s->tile_stage);
    s->tile_tmp = nullptr; s->tile_stage = nullptr; s->tile_tmp_bytes = s->tile_stage_bytes = 0;
    s->has_tmap = s->fused_ready = false;
    s->cur = 0; s->n = 0; s->cap = 0; s->n_entries = 0; s->arrived = 0; s->left = -1;
    s->counts_valid = s->perm_valid = s->moves_counted = s->sort_pending = false;
    for (int k = 0; k < 8; k++) s->epoch[k] = 0;
    s->step_no = 0; s->b2_pending = false; s->pending_ph = 3;
    s->kept_cap = 0;
    if (keep) { s->pa[0] = kp[0]; s->pa[1] = kp[1]; s->keys = kkeys; s->perm = kperm; s->dirs = kdirs; s->kept_cap = keep_cap; }
}
extern "C" {

int mpic_create(const mpic_params* p, mpic_sim** out) {
    REQUIRE(p && out, MPIC_EINVAL, "mpic_create: null argument");
    TRY(validate_params(p));
    const int gtr = p->ny / p->tny, gtc = p->nx / p->tnx;
    TRY(need_device());
    CUDA_TRY(cudaSetDevice(p->device));
    mpic_sim* s = new mpic_sim();
    s->prm = *p;
    TRY(mpic_set_geometry(s));
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, p->device));
    s->nsm = prop.multiProcessorCount;
    CUDA_TRY(cudaStreamCreateWithFlags(&s->st, cudaStreamNonBlocking));
    if (!getenv("MPIC_NO_SORT_OVERLAP")) {   // single slab: every step; slabs: the peer-ring step (mpic_peer_step_end)
        CUDA_TRY(cudaStreamCreateWithFlags(&s->st_sort, cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreateWithFlags(&s->ev_push, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&s->ev_sort, cudaEventDisableTiming));
    }
    TRY(mpic_alloc_state(s));
    s->owner.resize((size_t)gtr * gtc);
    for (int r = 0; r < gtr; r++) {
        int owner_rank = 0;
        for (int k = 0; k < p->nranks; k++) { int r0, nr; mpic_slab_rows(gtr, p->nranks, k, &r0, &nr); if (r >= r0 && r < r0 + nr) owner_rank = k; }
        if (p->nranks > 1 && r >= p->row0 && r < p->row0 + p->nrows) owner_rank = p->rank;
        for (int cidx = 0; cidx < gtc; cidx++) s->owner[(size_t)r * gtc + cidx] = owner_rank;
    }
    CUDA_TRY(cudaStreamSynchronize(s->st));
    *out = s;
    return 0;
}

int mpic_destroy(mpic_sim* s) {
    if (!s) return 0;
    cudaSetDevice(s->prm.device);
    if (s->st) cudaStreamSynchronize(s->st);
    dump_release(s);
    mpic_reslab_finish(s);
    mpic_free_state(s, 0);
    if (s->st_sort) { cudaStreamDestroy(s->st_sort); cudaEventDestroy(s->ev_push); cudaEventDestroy(s->ev_sort); }
    if (s->pipe_up) {
        cudaStreamDestroy(s->pipe_up); cudaStreamDestroy(s->pipe_down);
        for (int c = 0; c < MPIC_PIPE_MAX; c++) { cudaEventDestroy(s->pipe_ev_up[c]); cudaEventDestroy(s->pipe_ev_done[c]); }
        cudaEventDestroy(s->pipe_ev_free);
    }
    if (s->st) cudaStreamDestroy(s->st);
    delete s;
    return 0;
}
// device error word -> status (and reset).  Called wherever the host waits for the stream anyway.
static int check_error_word(mpic_sim* s) {
    unsigned int w = 0;
    CUDA_TRY(cudaMemcpyAsync(&w, s->G.err, 4, cudaMemcpyDeviceToHost, s->st));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    if (!w) return 0;
    CUDA_TRY(cudaMemsetAsync(s->G.err, 0, 4, s->st));
    REQUIRE(!(w & MPIC_ERR_POSITION), MPIC_EINVAL, "a particle position is outside [0,box_x) x [0,box_y) or not finite (binned into cell 0)");
    REQUIRE(!(w & MPIC_ERR_JRANGE), MPIC_EINVAL, "current accumulator out of range: |J| >= 1024 in code units exceeds the int64 fixed point (value * 2^52)");
    return 0;
}
int mpic_sync(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device); REQUIRE(s, MPIC_EINVAL, "null handle"); CUDA_TRY(cudaStreamSynchronize(s->st));
    TRY(check_error_word(s));
    if (s->peer_on) return mpic_peer_refresh_end(s);     // surfaces a timed-out neighbour wait
    return 0; }
void* mpic_stream(mpic_sim* s) { return s ? (void*)s->st : nullptr; }
int64_t mpic_kernel_launches(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device); return s ? s->launches : 0; }
int64_t mpic_num_particles(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device); return s ? s->n : 0; }

// ------------------------------------------------------------------------------------------ fields
static int refresh_all_guards(mpic_sim* s) {
    TRY(launch_wrap_guards(s->G, s->planes, 0, 6, s->G.ywrap, s->st));
    s->launches += s->G.ywrap ? 2 : 1;
    return 0;
}
}  // extern "C"
// grow-only device scratch (reused by the tile converters and occupancy queries: no cudaMalloc / cudaFree per call)
int mpic_scratch(void** buf, size_t* have, size_t need) {
    if (*have >= need && *buf) return 0;
    if (*buf) { cudaFree(*buf); *buf = nullptr; *have = 0; }
    const size_t want = need + need / 4 + 256;
    CUDA_TRY(cudaMalloc(buf, want));
    *have = want;
    return 0;
}
int mpic_refresh_x_guards(mpic_sim* s) { return refresh_all_guards(s); }
extern "C" {
int mpic_upload_planes(mpic_sim* s, const double* const planes[6]) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && planes, MPIC_EINVAL, "null argument");
    const FieldGeom& G = s->G;
    for (int p = 0; p < 6; p++)
        CUDA_TRY(cudaMemcpy2DAsync(s->planes + p * G.plane + (int64_t)G.g * G.pitch + MPIC_X0, (size_t)G.pitch * 8, planes[p],
                                   (size_t)G.nx * 8, (size_t)G.nx * 8, G.nyl, cudaMemcpyHostToDevice, s->st));
    TRY(refresh_all_guards(s));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    return 0;
}
int mpic_download_planes(mpic_sim* s, double* const planes[6]) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && planes, MPIC_EINVAL, "null argument");
    const FieldGeom& G = s->G;
    for (int p = 0; p < 6; p++)
        CUDA_TRY(cudaMemcpy2DAsync(planes[p], (size_t)G.nx * 8, s->planes + p * G.plane + (int64_t)G.g * G.pitch + MPIC_X0,
                                   (size_t)G.pitch * 8, (size_t)G.nx * 8, G.nyl, cudaMemcpyDeviceToHost, s->st));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    return 0;
}
static int tile_lists(mpic_sim* s, const int* gids, int ntiles, int** d_lists) {
    std::vector<int> h(2 * (size_t)ntiles);
    for (int t = 0; t < ntiles; t++) {
        int row = gids[t] / s->P.gtc, col = gids[t] % s->P.gtc;
        REQUIRE(gids[t] >= 0 && row >= s->prm.row0 && row < s->prm.row0 + s->prm.nrows, MPIC_EINVAL, "tile %d is not on this slab", gids[t]);
        h[t] = row - s->prm.row0; h[ntiles + t] = col;
    }
    TRY(mpic_scratch(&s->tile_tmp, &s->tile_tmp_bytes, h.size() * 4));
    *d_lists = (int*)s->tile_tmp;
    CUDA_TRY(cudaMemcpyAsync(*d_lists, h.data(), h.size() * 4, cudaMemcpyHostToDevice, s->st));
    CUDA_TRY(cudaStreamSynchronize(s->st));              // h goes out of scope
    return 0;
}
int mpic_upload_tiles(mpic_sim* s, const double* grid6, const int* gids, int ntiles) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && grid6 && gids && ntiles >= 0, MPIC_EINVAL, "bad argument");
    if (!ntiles) return 0;
    int* lists = nullptr;
    TRY(tile_lists(s, gids, ntiles, &lists));
    const size_t bytes = (size_t)ntiles * (s->P.tnx + 4) * (s->P.tny + 4) * 48;
    TRY(mpic_scratch(&s->tile_stage, &s->tile_stage_bytes, bytes));
    double* d = (double*)s->tile_stage;
    CUDA_TRY(cudaMemcpyAsync(d, grid6, bytes, cudaMemcpyHostToDevice, s->st));
    TRY(launch_tiles_to_planes(s->G, s->planes, d, lists, lists + ntiles, ntiles, s->P.tnx, s->P.tny, s->st));
    s->launches++;
    TRY(refresh_all_guards(s));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    return 0;
}
int mpic_download_tiles(mpic_sim* s, double* grid6, const int* gids, int ntiles) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && grid6 && gids && ntiles >= 0, MPIC_EINVAL, "bad argument");
    if (s->peer_on) TRY(mpic_sync(s));                   // guard rows are written by the neighbours: await their last phase
    if (!ntiles) return 0;
    int* lists = nullptr;
    TRY(tile_lists(s, gids, ntiles, &lists));
    const size_t bytes = (size_t)ntiles * (s->P.tnx + 4) * (s->P.tny + 4) * 48;
    TRY(mpic_scratch(&s->tile_stage, &s->tile_stage_bytes, bytes));
    double* d = (double*)s->tile_stage;
    TRY(launch_planes_to_tiles(s->G, s->planes, d, lists, lists + ntiles, ntiles, s->P.tnx, s->P.tny, s->st));
    s->launches++;
    CUDA_TRY(cudaMemcpyAsync(grid6, d, bytes, cudaMemcpyDeviceToHost, s->st));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    return 0;
}

int mpic_step_b_half(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s, MPIC_EINVAL, "null handle");
    TRY(launch_b_sweep(s->G, s->planes, s->cf, PeerHalo(), PeerSync(), s->st));
    s->launches++;
    return 0;
}
static int e_sweep(mpic_sim* s, int with_j) {
    TRY(launch_e_sweep(s->G, s->planes, s->jacc, with_j, s->cf, PeerHalo(), PeerSync(), s->st));
    s->launches++;
    return 0;
}
int mpic_step_e_full(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s, MPIC_EINVAL, "null handle");
    return e_sweep(s, s->cap > 0);      // J may have arrived from a neighbour slab even if this one holds no particle
}
int mpic_field_steps(mpic_sim* s, int nsteps) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && nsteps >= 0, MPIC_EINVAL, "bad argument");
    REQUIRE(s->prm.nranks == 1, MPIC_ESTATE, "mpic_field_steps needs a single slab; drive the phases + halos from the host for nranks>1");
    for (int k = 0; k < nsteps; k++) {
        TRY(mpic_step_b_half(s));
        TRY(e_sweep(s, 0));
        if (s->prm.scheme == 0) TRY(mpic_step_b_half(s));
    }
    return 0;
}
}  // extern "C"
int mpic_fused_setup(mpic_sim* s) {
    if (s->fused_ready) return 0;
    const size_t bytes = (size_t)6 * s->G.plane * 8;
    if (!s->planes_alt) {
        CUDA_TRY(cudaMalloc(&s->planes_alt, bytes));
        CUDA_TRY(cudaMemsetAsync(s->planes_alt, 0, bytes, s->st));
    }
    int pw, ph;
    fused_box_dims(&pw, &ph);
    TRY(encode_box_map(s, s->planes, pw, ph, &s->ftmap));
    TRY(encode_box_map(s, s->planes_alt, pw, ph, &s->ftmap_alt));
    if (s->has_tmap) {
        push_box_dims(s->P.tnx, s->P.tny, &pw, &ph);
        TRY(encode_box_map(s, s->planes_alt, pw, ph, &s->tmap_alt));
        pack_box_dims(s->P.tnx, s->P.tny, &pw, &ph);
        TRY(encode_box_map(s, s->planes_alt, pw, ph, &s->tmap_pack_alt));
    }
    s->fused_ready = true;
    return 0;
}
void mpic_fused_swap(mpic_sim* s) {
    std::swap(s->planes, s->planes_alt);
    std::swap(s->ftmap, s->ftmap_alt);
    if (s->has_tmap) { std::swap(s->tmap, s->tmap_alt); std::swap(s->tmap_pack, s->tmap_pack_alt); }
    for (int side = 0; side < 2; side++) std::swap(s->peer[side].planes, s->peer[side].planes_alt);
}
// one fused half-B / E / half-B step of a single slab (periodic in y inside the kernel's images)
static int fused_step_local(mpic_sim* s, int with_j) {
    TRY(mpic_fused_setup(s));
    TRY(launch_fused_step(s->G, &s->ftmap, s->planes_alt, s->jacc, with_j, s->cf, PeerHalo(), PeerSync(), s->st));
    s->launches++;
    if (with_j) CUDA_TRY(cudaMemsetAsync(s->jacc, 0, (size_t)3 * s->G.plane * 8, s->st));   // the accumulators are consumed
    mpic_fused_swap(s);
    return 0;
}
extern "C" {
int mpic_step_fields_fused(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s, MPIC_EINVAL, "null handle");
    REQUIRE(s->prm.nranks == 1 && s->prm.scheme == 0, MPIC_ESTATE, "mpic_step_fields_fused: single slab, scheme 0 only");
    return fused_step_local(s, s->cap > 0);
}
int mpic_field_steps_fused(mpic_sim* s, int nsteps) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && nsteps >= 0, MPIC_EINVAL, "bad argument");
    REQUIRE(s->prm.nranks == 1, MPIC_ESTATE, "mpic_field_steps_fused needs a single slab (slabs: mpic_peer_field_step_fused)");
    if (s->prm.scheme != 0) return mpic_field_steps(s, nsteps);      // v1 scheme (B full dt, E): two sweeps, nothing to fuse across
    for (int k = 0; k < nsteps; k++) TRY(fused_step_local(s, 0));
    return 0;
}

// ------------------------------------------------------------------------------------------ halos
static double* halo_base(mpic_sim* s, int which) {
    return which == 2 ? (double*)s->jacc : s->planes + (which == 1 ? 3 * s->G.plane : 0);
}
int mpic_halo_describe(mpic_sim* s, mpic_halo_desc* out) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && out, MPIC_EINVAL, "bad argument");
    REQUIRE(s->halo_send, MPIC_ESTATE, "halos exist only for nranks > 1");
    for (int d = 0; d < 2; d++) { out->send[d] = s->halo_send + d * s->halo_side; out->recv[d] = s->halo_recv + d * s->halo_side; }
    out->bytes = s->halo_side * 8;
    return 0;
}
__global__ void halo_header_kernel(double* send, int64_t side_stride, const int* leave_count) {
    if (threadIdx.x < 2) ((long long*)(send + threadIdx.x * side_stride))[0] = leave_count[threadIdx.x];
}
int mpic_halo_pack(mpic_sim* s, int which) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && which >= 0 && which <= 2, MPIC_EINVAL, "bad argument");
    REQUIRE(s->halo_send, MPIC_ESTATE, "halos exist only for nranks > 1");
    TRY(launch_halo_pack(s->G, halo_base(s, which), which == 2, s->halo_send + HALO_HDR, s->halo_side, s->st));
    s->launches++;
    if (which == 2) { halo_header_kernel<<<1, 32, 0, s->st>>>(s->halo_send, s->halo_side, s->lv.count); s->launches++; }
    return 0;
}
int mpic_halo_unpack(mpic_sim* s, int which) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && which >= 0 && which <= 2, MPIC_EINVAL, "bad argument");
    REQUIRE(s->halo_recv, MPIC_ESTATE, "halos exist only for nranks > 1");
    TRY(launch_halo_unpack(s->G, halo_base(s, which), which == 2, s->halo_recv + HALO_HDR, s->halo_side, s->st));
    s->launches++;
    return 0;
}

// ------------------------------------------------------------------------------------------ particle migration between slabs
int mpic_migration_describe(mpic_sim* s, mpic_migr_desc* out) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && out, MPIC_EINVAL, "bad argument");
    REQUIRE(s->lv.rec[0], MPIC_ESTATE, "migration buffers exist only for nranks > 1 with particle storage");
    for (int d = 0; d < 2; d++) { out->send[d] = s->lv.rec[d]; out->recv[d] = s->arrive[d]; }
    out->cap = s->lv.cap; out->record_bytes = MPIC_REC_DOUBLES * 8;
    return 0;
}
int mpic_migration_counts(mpic_sim* s, int leave[2], int arrive[2]) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && leave && arrive, MPIC_EINVAL, "bad argument");
    REQUIRE(s->lv.rec[0], MPIC_ESTATE, "migration buffers exist only for nranks > 1 with particle storage");
    long long hdr[2];
    CUDA_TRY(cudaMemcpyAsync(leave, s->lv.count, 2 * sizeof(int), cudaMemcpyDeviceToHost, s->st));
    // the neighbours' leaver counts travelled in the header of the current (which=2) halo message
    CUDA_TRY(cudaMemcpyAsync(&hdr[0], s->halo_recv, 8, cudaMemcpyDeviceToHost, s->st));
    CUDA_TRY(cudaMemcpyAsync(&hdr[1], s->halo_recv + s->halo_side, 8, cudaMemcpyDeviceToHost, s->st));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    arrive[0] = (int)hdr[0]; arrive[1] = (int)hdr[1];
    for (int d = 0; d < 2; d++) {
        REQUIRE(leave[d] <= s->lv.cap && arrive[d] <= s->lv.cap, MPIC_ENOMEM, "migration buffer overflow: %d leavers / %d arrivals > capacity %d",
                leave[d], arrive[d], s->lv.cap);
    }
    s->left = (int64_t)leave[0] + leave[1];
    s->migrated_total += s->left;
    return 0;
}
int64_t mpic_migrated_total(mpic_sim* s) { return s ? s->migrated_total : 0; }
int mpic_migration_absorb(mpic_sim* s, int n_from_lo, int n_from_hi) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && n_from_lo >= 0 && n_from_hi >= 0, MPIC_EINVAL, "bad argument");
    REQUIRE(s->lv.rec[0], MPIC_ESTATE, "migration buffers exist only for nranks > 1 with particle storage");
    REQUIRE(s->n + s->arrived + n_from_lo + n_from_hi <= s->cap, MPIC_ENOMEM, "particle capacity exceeded by arrivals");
    REQUIRE(s->counts_valid, MPIC_ESTATE, "mpic_migration_absorb must follow mpic_push_deposit");
    // arrivals are appended in (source rank order: lower, upper; source order) -- SURVEY.md B.8
    TRY(launch_absorb(s->P, s->pa[s->cur], s->n + s->arrived, s->arrive[0], n_from_lo, s->cell_count, s->keys, s->G.err, s->st));
    s->arrived += n_from_lo;
    TRY(launch_absorb(s->P, s->pa[s->cur], s->n + s->arrived, s->arrive[1], n_from_hi, s->cell_count, s->keys, s->G.err, s->st));
    s->arrived += n_from_hi;
    s->launches += (n_from_lo > 0) + (n_from_hi > 0);
    return 0;
}

// ------------------------------------------------------------------------------------------ particles
static int sort_all(mpic_sim* s, bool counted, cudaStream_t force = nullptr);
// The pipelined host-buffer step leaves the resident particles chunk by chunk and defers the tile sort to the first call that
// needs sorted resident state (a loop of host-buffer steps overwrites that state every call and never pays for the sort).
static int ensure_sorted(mpic_sim* s) {
    if (!s->sort_pending) return 0;
    s->sort_pending = false;
    return sort_all(s, false, nullptr);
}
// physical reorder on demand: after this pa[cur][0..n) is itself in sorted order
static int materialise(mpic_sim* s) {
    TRY(ensure_sorted(s));
    if (!s->perm_valid) return 0;
    TRY(launch_gather(s->pa[s->cur], s->pa[s->cur ^ 1], s->perm, s->n, nullptr, s->st));
    s->launches++;
    s->cur ^= 1;
    s->perm_valid = false;
    return 0;
}
// The sort (kernels in particles.cu).  Both forms are STABLE and reproducible: no atomic decides a final slot.
// counted (after a push): the push left every particle's direction code in dirs[] and the per-cell move counts in mcnt, and
//   mpic_push_deposit turned them into the destination histogram cell_count (arrivals absorbed since were added to it):
//   scan -> offsets -> perm_stable (+ arrivals appended); no particle data moves, the next push gathers through perm.
// otherwise (uploads, loads, appends: arbitrary storage order): keys from the stored positions, histogram, scan, provisional
//   permutation, per-cell ascending sort of the permutation (== stable order), physical gather.
static int sort_all(mpic_sim* s, bool counted, cudaStream_t force) {
    TRY(ensure_sorted(s));
    counted = counted && s->counts_valid;
    s->counts_valid = false;
    const int64_t total = s->n + s->arrived;
    if (counted) {
        // single slab: on the sort stream, behind the push only -- whatever was queued on the main stream since (the field
        // update: it touches the planes and the current, the sort the move counts and the permutation) runs next to it; the
        // main stream continues once both are done
        // (peer-ring slabs: mpic_peer_step_end forces the sort stream -- already ordered behind the migration counts -- so that
        // the absorb of the arrivals and the sort run next to the E and second half-B sweeps)
        cudaStream_t st = s->st;
        if (force) st = force;
        else if (s->st_sort && s->prm.nranks == 1 && s->arrived == 0) { st = s->st_sort; CUDA_TRY(cudaStreamWaitEvent(st, s->ev_push, 0)); }
        if (!s->moves_counted) { TRY(launch_count_moves(s->P, s->G.ywrap, s->mcnt, s->cell_count, st)); s->launches++; }
        s->moves_counted = false;
        TRY(launch_scan_cells(s->cell_count, s->cell_start_alt, s->n_entries, s->scan_tmp, st));
        TRY(launch_offsets(s->P, s->G.ywrap, s->mcnt, s->cell_start_alt, s->offtab, s->cell_count, st));
        TRY(launch_perm_stable(s->P, s->G.ywrap, s->cell_start, s->dirs, s->offtab, s->perm, s->n, st));
        TRY(launch_arrivals(s->P, s->keys, s->n, (int)s->arrived, s->cell_count, s->cell_start_alt, s->perm, st));
        if (st != s->st) { CUDA_TRY(cudaEventRecord(s->ev_sort, st)); CUDA_TRY(cudaStreamWaitEvent(s->st, s->ev_sort, 0)); }
        std::swap(s->cell_start, s->cell_start_alt);
        s->perm_valid = true;
        s->launches += 5 + (s->arrived > 0 ? 2 : 0);
    } else {
        if (s->perm_valid && s->arrived == 0) return 0;          // already sorted, nothing new
        TRY(materialise(s));
        CUDA_TRY(cudaMemsetAsync(s->cell_count, 0, (size_t)s->n_entries * 4, s->st));
        TRY(launch_count_keys(s->P, s->pa[s->cur], 0, total, s->cell_count, s->keys, s->G.err, s->st));
        TRY(launch_scan_cells(s->cell_count, s->cell_start, s->n_entries, s->scan_tmp, s->st));
        TRY(launch_perm_build(s->P, s->keys, total, s->cell_count, s->perm, s->st));
        TRY(launch_segsort(s->cell_start, s->P.ncells, s->perm, s->keys, s->st));
        // physical reorder; slab leavers (keys >= ncells) are not in perm: only the first cell_start[ncells] slots exist
        TRY(launch_gather(s->pa[s->cur], s->pa[s->cur ^ 1], s->perm, total, s->cell_start + s->P.ncells, s->st));
        s->cur ^= 1;
        s->perm_valid = false;
        s->launches += 8;
    }
    if (s->prm.nranks > 1) {       // particles whose key is a leaver bin are not entered in perm
        if (counted && s->left >= 0) s->n = total - s->left;
        else {
            int kept = 0;
            CUDA_TRY(cudaMemcpyAsync(&kept, s->cell_start + s->P.ncells, sizeof(int), cudaMemcpyDeviceToHost, s->st));
            CUDA_TRY(cudaStreamSynchronize(s->st));
            s->n = kept;
        }
    }
    s->arrived = 0; s->left = -1;
    return 0;
}
int mpic_upload_particles(mpic_sim* s, int64_t n, const double* x, const double* y, const double* ux, const double* uy,
                          const double* uz, const double* q, const int* tag) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && n >= 0, MPIC_EINVAL, "bad argument");
    REQUIRE(n == 0 || (x && y && ux && uy && uz && q && tag), MPIC_EINVAL, "mpic_upload_particles: null particle array");
    REQUIRE(n <= s->cap, MPIC_ENOMEM, "particle capacity %lld exceeded (%lld)", (long long)s->cap, (long long)n);
    s->perm_valid = false; s->counts_valid = false; s->arrived = 0; s->sort_pending = false;
    Particles& p = s->pa[s->cur];
    const double* src[6] = { x, y, ux, uy, uz, q };
    double* dst[6] = { p.x, p.y, p.ux, p.uy, p.uz, p.q };
    for (int k = 0; k < 6; k++) CUDA_TRY(cudaMemcpyAsync(dst[k], src[k], (size_t)n * 8, cudaMemcpyHostToDevice, s->st));
    CUDA_TRY(cudaMemcpyAsync(p.tag, tag, (size_t)n * 4, cudaMemcpyHostToDevice, s->st));
    s->n = n;
    TRY(sort_all(s, false));
    return check_error_word(s);        // positions outside the box are reported here, not at some later launch
}
int mpic_download_particles(mpic_sim* s, int64_t cap, double* x, double* y, double* ux, double* uy, double* uz, double* q,
                            int* tag, int64_t* n_out) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s, MPIC_EINVAL, "null handle");
    REQUIRE(cap >= s->n, MPIC_ENOMEM, "host buffers too small");
    TRY(materialise(s));
    Particles& p = s->pa[s->cur];
    double* dst[6] = { x, y, ux, uy, uz, q };
    const double* src[6] = { p.x, p.y, p.ux, p.uy, p.uz, p.q };
    for (int k = 0; k < 6; k++) if (dst[k]) CUDA_TRY(cudaMemcpyAsync(dst[k], src[k], (size_t)s->n * 8, cudaMemcpyDeviceToHost, s->st));
    if (tag) CUDA_TRY(cudaMemcpyAsync(tag, p.tag, (size_t)s->n * 4, cudaMemcpyDeviceToHost, s->st));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    if (n_out) *n_out = s->n;
    return 0;
}
int mpic_load_uniform(mpic_sim* s, int ppc, double u_th, double d0, double d1, double d2, double two_stream, double density,
                      double band_lo, double band_hi, uint64_t seed, int append) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && ppc > 0 && s->cap > 0, MPIC_EINVAL, "bad argument");
    const PicConst& P = s->P;
    int lo = (int)nearbyint(band_lo * P.ny), hi = (int)nearbyint(band_hi * P.ny);
    REQUIRE(lo >= 0 && hi <= P.ny && lo <= hi, MPIC_EINVAL, "bad band");
    const int64_t tag0 = append ? s->tag_base : 0;
    if (!append) { s->n = 0; s->tag_base = 0; s->perm_valid = false; s->sort_pending = false; }
    TRY(materialise(s));
    s->counts_valid = false;
    const int slo = lo > P.cy0 ? lo : P.cy0, shi = hi < P.cy0 + P.nyl ? hi : P.cy0 + P.nyl;
    const int64_t nsel = shi > slo ? (int64_t)(shi - slo) * P.nx : 0;
    REQUIRE(s->n + nsel * ppc <= s->cap, MPIC_ENOMEM, "particle capacity exceeded");
    const double qp = -density * P.dx * P.dy / ppc;
    TRY(launch_load_uniform(P, s->pa[s->cur], s->n, nsel, slo, tag0, ppc, u_th, d0, d1, d2, two_stream, qp, seed, s->st));
    s->launches++;
    s->n += nsel * ppc;
    s->tag_base = tag0 + (int64_t)(hi - lo) * P.nx * ppc;
    TRY(sort_all(s, false));
    return 0;
}

int mpic_push_deposit(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && s->cap > 0, MPIC_ESTATE, "no particle storage");
    TRY(ensure_sorted(s));
    REQUIRE(s->arrived == 0, MPIC_ESTATE, "arrivals absorbed but not sorted in: call mpic_sort_tiles before the next push");
    REQUIRE(!s->counts_valid, MPIC_ESTATE, "particles moved but not re-sorted: call mpic_sort_tiles between pushes");
    s->left = s->prm.nranks > 1 ? -1 : 0;
    TRY(launch_push_deposit(s->P, s->G, &s->tmap, &s->tmap_pack, s->jacc, s->pa[s->cur], s->pa[s->cur ^ 1], s->perm,
                            s->cell_start, s->mcnt, s->dirs, s->push_split, s->lv, s->st, s->n));
    s->cur ^= 1;                    // the push wrote the other buffer in slot order
    s->perm_valid = false;
    // destination histogram of the push (a gather over the move counts, no atomics); slab leavers: their records are written in
    // slot order into the send buffers, their counts into lv.count and the two leaver bins of the histogram
    s->launches++;
    s->moves_counted = false;
    // single slab, and the peer-ring step on slabs: the histogram is formed by the sort, on its stream
    if (s->st_sort && (s->prm.nranks == 1 || s->defer_count)) CUDA_TRY(cudaEventRecord(s->ev_push, s->st));
    else { TRY(launch_count_moves(s->P, s->G.ywrap, s->mcnt, s->cell_count, s->st)); s->launches++; s->moves_counted = true; }
    if (s->prm.nranks > 1 && !(s->st_sort && s->defer_count)) {     // (peer-ring step: queued on the side stream by mpic_peer_step_begin)
        TRY(launch_leavers(s->P, s->pa[s->cur], s->cell_start, s->dirs, s->mcnt, s->lv_off, s->cell_count, s->lv, s->st));
        s->launches += 2;
    }
    s->counts_valid = true;
    return 0;
}
int mpic_sort_after_push(mpic_sim* s) { return sort_all(s, true); }
int mpic_sort_after_push_on(mpic_sim* s, void* stream) { return sort_all(s, true, (cudaStream_t)stream); }
int mpic_sort_uploaded(mpic_sim* s) { s->perm_valid = false; s->counts_valid = false; s->arrived = 0; s->sort_pending = false; return sort_all(s, false); }
int mpic_ensure_sorted(mpic_sim* s) { return ensure_sorted(s); }
int mpic_sort_tiles(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && s->cap > 0, MPIC_ESTATE, "no particle storage");
    return sort_all(s, true);      // reuses the histogram the push produced when there is one
}
static int fold_current(mpic_sim* s) {
    TRY(launch_fold_current(s->G, s->jacc, 3, s->G.ywrap, s->st));
    s->launches += s->G.ywrap ? 2 : 1;
    return 0;
}
int mpic_pic_steps(mpic_sim* s, int nsteps) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && nsteps >= 0, MPIC_EINVAL, "bad argument");
    REQUIRE(s->cap > 0, MPIC_ESTATE, "no particle storage");
    REQUIRE(s->prm.nranks == 1, MPIC_ESTATE, "mpic_pic_steps needs a single slab; drive phases + halos from the host for nranks>1");
    static const bool fuse = !getenv("MPIC_NO_FUSED_YEE");
    for (int k = 0; k < nsteps; k++) {
        TRY(mpic_push_deposit(s));
        if (fuse && s->prm.scheme == 0) TRY(fused_step_local(s, 1));   // half-B, E - dt*J (guard ring folded on the fly), half-B: one kernel
        else {
            TRY(mpic_step_b_half(s));                   // the E sweep folds the J guard ring on the fly (yee.cu)
            TRY(e_sweep(s, 1));
            TRY(mpic_step_b_half(s));
        }
        TRY(sort_all(s, true));
    }
    return 0;
}
static Particles offset_particles(const Particles& p, int64_t o) {
    Particles r = { p.x + o, p.y + o, p.ux + o, p.uy + o, p.uz + o, p.q + o, p.tag + o };
    return r;
}
static int pipe_setup(mpic_sim* s) {
    if (!s->pipe_count) {
        CUDA_TRY(cudaMalloc(&s->pipe_count, (size_t)s->n_entries * 4));
        CUDA_TRY(cudaMalloc(&s->pipe_start, (size_t)(s->n_entries + 1) * 4));
    }
    if (s->pipe_up) return 0;
    CUDA_TRY(cudaStreamCreateWithFlags(&s->pipe_up, cudaStreamNonBlocking));
    CUDA_TRY(cudaStreamCreateWithFlags(&s->pipe_down, cudaStreamNonBlocking));
    for (int c = 0; c < MPIC_PIPE_MAX; c++) {
        CUDA_TRY(cudaEventCreateWithFlags(&s->pipe_ev_up[c], cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&s->pipe_ev_done[c], cudaEventDisableTiming));
    }
    CUDA_TRY(cudaEventCreateWithFlags(&s->pipe_ev_free, cudaEventDisableTiming));
    return 0;
}

// One PIC step on HOST buffers with the PCIe transfers overlapped with the compute (single slab, one step): the particles
// cross the bus in chunks; while chunk c+1 is still uploading, chunk c is sorted (count / scan / scatter inside its own
// index range), pushed (the tile CTAs of tiles without particles in this chunk exit at once) and its results are already
// on their way back on a second copy stream -- the two directions of the link run concurrently.  The current of all chunks
// accumulates in the same int64 planes (integer adds: any order, same bits), the Yee sweeps follow the last chunk, and the
// tile sort of the resident state is the usual scan + permutation build over the keys every push wrote.  Same results as
// upload -> mpic_pic_steps(1) -> download; the particles come back in a different ORDER (chunk by chunk, tile-sorted
// inside each chunk), which the API never promised.
static int pic_step_host_pipelined(mpic_sim* s, double* const planes[6], int64_t n, double* x, double* y, double* ux,
                                   double* uy, double* uz, double* q, int* tag, int64_t* n_out) {
    TRY(pipe_setup(s));
    const FieldGeom& G = s->G;
    int64_t chunk = 1 << 20;                                       // ~1 M particles (52 MB) per chunk (B200: 21.0 ms; 2 M: 22.3; 4 M: 23.2)
    if (const char* e = getenv("MPIC_PIPE_CHUNK")) { const long long v = atoll(e); if (v >= 1024) chunk = v; }   // tuning knob
    int nchunk = (int)((n + chunk - 1) / chunk);
    if (nchunk < 2) nchunk = 2;
    if (nchunk > MPIC_PIPE_MAX) nchunk = MPIC_PIPE_MAX;
    s->perm_valid = false; s->counts_valid = false; s->arrived = 0; s->sort_pending = false;   // the resident particles are replaced
    Particles& A = s->pa[s->cur];
    Particles& B = s->pa[s->cur ^ 1];
    // everything enqueued earlier on the main stream (it may still read A/B) precedes the first upload
    CUDA_TRY(cudaEventRecord(s->pipe_ev_free, s->st));
    CUDA_TRY(cudaStreamWaitEvent(s->pipe_up, s->pipe_ev_free, 0));
    CUDA_TRY(cudaStreamWaitEvent(s->pipe_down, s->pipe_ev_free, 0));
    for (int p = 0; p < 6; p++)
        CUDA_TRY(cudaMemcpy2DAsync(s->planes + p * G.plane + (int64_t)G.g * G.pitch + MPIC_X0, (size_t)G.pitch * 8, planes[p],
                                   (size_t)G.nx * 8, (size_t)G.nx * 8, G.nyl, cudaMemcpyHostToDevice, s->st));
    TRY(refresh_all_guards(s));
    const double* src[6] = { x, y, ux, uy, uz, q };
    double* hdst[6] = { x, y, ux, uy, uz, q };
    for (int c = 0; c < nchunk; c++) {
        const int64_t lo = n * c / nchunk, hi = n * (c + 1) / nchunk, m = hi - lo;
        if (m <= 0) continue;
        double* dA[6] = { A.x, A.y, A.ux, A.uy, A.uz, A.q };
        for (int k = 0; k < 6; k++) CUDA_TRY(cudaMemcpyAsync(dA[k] + lo, src[k] + lo, (size_t)m * 8, cudaMemcpyHostToDevice, s->pipe_up));
        CUDA_TRY(cudaMemcpyAsync(A.tag + lo, tag + lo, (size_t)m * 4, cudaMemcpyHostToDevice, s->pipe_up));
        CUDA_TRY(cudaEventRecord(s->pipe_ev_up[c], s->pipe_up));
        // sort the chunk inside [lo, hi) (stable: provisional permutation + per-cell ascending sort), then push A -> B through it
        CUDA_TRY(cudaStreamWaitEvent(s->st, s->pipe_ev_up[c], 0));
        CUDA_TRY(cudaMemsetAsync(s->pipe_count, 0, (size_t)s->n_entries * 4, s->st));
        const Particles Ac = offset_particles(A, lo), Bc = offset_particles(B, lo);
        TRY(launch_count_keys(s->P, Ac, 0, m, s->pipe_count, s->keys + lo, s->G.err, s->st));
        TRY(launch_scan_cells(s->pipe_count, s->pipe_start, s->n_entries, s->scan_tmp, s->st));
        TRY(launch_perm_build(s->P, s->keys + lo, m, s->pipe_count, s->perm + lo, s->st));
        TRY(launch_segsort(s->pipe_start, s->P.ncells, s->perm + lo, s->keys + lo, s->st));
        TRY(launch_push_deposit(s->P, s->G, &s->tmap, &s->tmap_pack, s->jacc, Ac, Bc, s->perm + lo, s->pipe_start, s->mcnt, s->dirs + lo,
                                s->push_split, s->lv, s->st, 0));      // chunk occupancy unknown: adaptive lane groups
        s->launches += 8;
        CUDA_TRY(cudaEventRecord(s->pipe_ev_done[c], s->st));
        // results of the chunk go home while the next one is still arriving
        CUDA_TRY(cudaStreamWaitEvent(s->pipe_down, s->pipe_ev_done[c], 0));
        double* dB[6] = { B.x, B.y, B.ux, B.uy, B.uz, B.q };
        for (int k = 0; k < 6; k++) CUDA_TRY(cudaMemcpyAsync(hdst[k] + lo, dB[k] + lo, (size_t)m * 8, cudaMemcpyDeviceToHost, s->pipe_down));
        CUDA_TRY(cudaMemcpyAsync(tag + lo, B.tag + lo, (size_t)m * 4, cudaMemcpyDeviceToHost, s->pipe_down));
    }
    s->cur ^= 1;                       // the pushed particles are in B, chunk by chunk
    s->n = n;
    s->left = 0;
    if (s->prm.scheme == 0 && !getenv("MPIC_NO_FUSED_YEE")) TRY(fused_step_local(s, 1));
    else {
        TRY(mpic_step_b_half(s));
        TRY(e_sweep(s, 1));
        TRY(mpic_step_b_half(s));
    }
    for (int p = 0; p < 6; p++)
        CUDA_TRY(cudaMemcpy2DAsync(planes[p], (size_t)G.nx * 8, s->planes + p * G.plane + (int64_t)G.g * G.pitch + MPIC_X0,
                                   (size_t)G.pitch * 8, (size_t)G.nx * 8, G.nyl, cudaMemcpyDeviceToHost, s->st));
    s->sort_pending = true;            // resident state stays usable: one stable sort by position over all chunks, run by the first
                                       // call that needs it (ensure_sorted) -- a loop of host-buffer steps never does
    CUDA_TRY(cudaStreamSynchronize(s->pipe_down));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    if (n_out) *n_out = n;
    return check_error_word(s);
}

int mpic_pic_steps_host(mpic_sim* s, int nsteps, double* const planes[6], int64_t n, double* x, double* y, double* ux,
                        double* uy, double* uz, double* q, int* tag, int64_t* n_out) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && planes, MPIC_EINVAL, "bad argument");
    REQUIRE(n >= 0 && n <= s->cap, MPIC_ENOMEM, "particle capacity %lld exceeded (%lld)", (long long)s->cap, (long long)n);
    REQUIRE(n == 0 || (x && y && ux && uy && uz && q && tag), MPIC_EINVAL, "mpic_pic_steps_host: null particle array");
    for (int p = 0; p < 6; p++) REQUIRE(planes[p], MPIC_EINVAL, "mpic_pic_steps_host: null field plane %d", p);
    int64_t pipe_min = 1 << 20;        // below this the transfers are too short to be worth chunking
    if (const char* e = getenv("MPIC_PIPELINE_MIN")) pipe_min = atoll(e);
    if (nsteps == 1 && s->prm.nranks == 1 && n >= pipe_min && n >= 2 && !getenv("MPIC_NO_PIPELINE"))
        return pic_step_host_pipelined(s, planes, n, x, y, ux, uy, uz, q, tag, n_out);
    const FieldGeom& G = s->G;
    for (int p = 0; p < 6; p++)
        CUDA_TRY(cudaMemcpy2DAsync(s->planes + p * G.plane + (int64_t)G.g * G.pitch + MPIC_X0, (size_t)G.pitch * 8, planes[p],
                                   (size_t)G.nx * 8, (size_t)G.nx * 8, G.nyl, cudaMemcpyHostToDevice, s->st));
    TRY(refresh_all_guards(s));
    TRY(mpic_upload_particles(s, n, x, y, ux, uy, uz, q, tag));
    TRY(mpic_pic_steps(s, nsteps));
    for (int p = 0; p < 6; p++)
        CUDA_TRY(cudaMemcpy2DAsync(planes[p], (size_t)G.nx * 8, s->planes + p * G.plane + (int64_t)G.g * G.pitch + MPIC_X0,
                                   (size_t)G.pitch * 8, (size_t)G.nx * 8, G.nyl, cudaMemcpyDeviceToHost, s->st));
    return mpic_download_particles(s, n, x, y, ux, uy, uz, q, tag, n_out);
}

int mpic_cell_crossers(mpic_sim* s, int64_t* n_out) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && n_out && s->cap > 0, MPIC_EINVAL, "bad argument");
    REQUIRE(s->counts_valid && s->arrived == 0, MPIC_ESTATE, "mpic_cell_crossers: call between mpic_push_deposit and mpic_sort_tiles");
    CUDA_TRY(cudaMemsetAsync(s->d_scratch, 0, 8, s->st));
    TRY(launch_crossers(s->dirs, s->n, (unsigned long long*)s->d_scratch, s->st));
    s->launches++;
    unsigned long long h = 0;
    CUDA_TRY(cudaMemcpyAsync(&h, s->d_scratch, 8, cudaMemcpyDeviceToHost, s->st));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    *n_out = (int64_t)h;
    return 0;
}
int mpic_get_tile_counts(mpic_sim* s, int* counts) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && counts && s->cap > 0, MPIC_EINVAL, "bad argument");
    TRY(ensure_sorted(s));
    const int nt = s->P.ltr * s->P.gtc;
    TRY(mpic_scratch(&s->tile_tmp, &s->tile_tmp_bytes, (size_t)nt * 4));
    int* d = (int*)s->tile_tmp;
    TRY(launch_tile_counts(s->P, s->cell_start, d, s->st));
    s->launches++;
    CUDA_TRY(cudaMemcpyAsync(counts, d, (size_t)nt * 4, cudaMemcpyDeviceToHost, s->st));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    return 0;
}
int mpic_get_particle_tile_index(mpic_sim* s, int* gid) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && gid && s->cap > 0, MPIC_EINVAL, "bad argument");
    if (!s->n) return 0;
    TRY(materialise(s));
    TRY(mpic_scratch(&s->tile_tmp, &s->tile_tmp_bytes, (size_t)s->n * 4));
    int* d = (int*)s->tile_tmp;
    TRY(launch_particle_tiles(s->P, s->pa[s->cur], s->n, s->prm.row0, d, s->st));
    s->launches++;
    CUDA_TRY(cudaMemcpyAsync(gid, d, (size_t)s->n * 4, cudaMemcpyDeviceToHost, s->st));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    return 0;
}
int mpic_download_current(mpic_sim* s, int64_t* const jplanes[3]) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && jplanes, MPIC_EINVAL, "bad argument");
    const FieldGeom& G = s->G;
    for (int p = 0; p < 3; p++)
        CUDA_TRY(cudaMemcpy2DAsync(jplanes[p], (size_t)G.nx * 8, s->jacc + p * G.plane + (int64_t)G.g * G.pitch + MPIC_X0,
                                   (size_t)G.pitch * 8, (size_t)G.nx * 8, G.nyl, cudaMemcpyDeviceToHost, s->st));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    return 0;
}
int mpic_fold_current(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device); REQUIRE(s, MPIC_EINVAL, "null handle"); return fold_current(s); }

// rho of the stored particles (node weights) into s->rho, guard ring folded as far as this slab can
static int deposit_rho_plane(mpic_sim* s) {
    const FieldGeom& G = s->G;
    TRY(materialise(s));
    if (!s->rho) CUDA_TRY(cudaMalloc(&s->rho, (size_t)G.plane * 8));
    CUDA_TRY(cudaMemsetAsync(s->rho, 0, (size_t)G.plane * 8, s->st));
    TRY(launch_deposit_rho(s->P, G, s->pa[s->cur], s->n, s->rho, s->st));
    TRY(launch_fold_current(G, s->rho, 1, G.ywrap, s->st));
    s->launches += 3;
    return 0;
}
int mpic_download_rho(mpic_sim* s, int64_t* rho) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && rho && s->cap > 0, MPIC_EINVAL, "bad argument");
    const FieldGeom& G = s->G;
    TRY(deposit_rho_plane(s));
    CUDA_TRY(cudaMemcpy2DAsync(rho, (size_t)G.nx * 8, s->rho + (int64_t)G.g * G.pitch + MPIC_X0, (size_t)G.pitch * 8, (size_t)G.nx * 8,
                               G.nyl, cudaMemcpyDeviceToHost, s->st));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    return 0;
}
int mpic_diagnostics(mpic_sim* s, double out[5]) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && out, MPIC_EINVAL, "bad argument");
    const FieldGeom& G = s->G;
    if (s->peer_on) TRY(mpic_sync(s));
    CUDA_TRY(cudaMemsetAsync(s->d_scratch, 0, 64, s->st));
    TRY(launch_field_energy(G, s->planes, s->d_scratch, s->d_scratch + 16, s->st));
    s->launches++;
    double h[8] = { 0 };
    double mean_rho = 0.0;
    if (s->cap > 0 && s->n > 0) {
        TRY(materialise(s));
        TRY(launch_kinetic(s->pa[s->cur], s->n, s->d_scratch + 1, s->d_scratch + 16, s->st));
        s->launches++;
        TRY(deposit_rho_plane(s));
    }
    CUDA_TRY(cudaMemcpyAsync(h, s->d_scratch, 64, cudaMemcpyDeviceToHost, s->st));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    out[0] = 0.5 * h[0] * s->P.dx * s->P.dy;
    out[1] = h[1] / fabs(s->prm.qm);
    out[3] = h[2];
    out[4] = (double)s->n;
    out[2] = 0.0;
    if (s->rho && s->n > 0) {
        // single slab: the neutralising background is minus the mean charge density.  Slabs: the LOCAL mean is used and the
        // first row is skipped (see mpic_gauss_residual, which takes the global mean)
        mean_rho = h[2] * s->P.cjz / ((double)G.nx * G.nyl);          // sum q /(dx dy) / cells
        TRY(launch_gauss_residual(G, s->planes, s->rho, mean_rho, s->P.inv_dx, s->P.inv_dy, s->d_scratch + 4, G.ywrap ? 0 : 1, s->st));
        s->launches++;
        CUDA_TRY(cudaMemcpyAsync(h, s->d_scratch, 64, cudaMemcpyDeviceToHost, s->st));
        CUDA_TRY(cudaStreamSynchronize(s->st));
        out[2] = h[4];
    }
    return 0;
}
int mpic_gauss_residual(mpic_sim* s, double mean_rho, double* out) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && out, MPIC_EINVAL, "bad argument");
    REQUIRE(s->cap > 0, MPIC_ESTATE, "no particle storage");
    if (s->peer_on) TRY(mpic_sync(s));
    TRY(deposit_rho_plane(s));
    CUDA_TRY(cudaMemsetAsync(s->d_scratch, 0, 64, s->st));
    // a slab's row 0 misses the node charge of the particles in the lower neighbour's top cell row: rows 1.. are complete
    TRY(launch_gauss_residual(s->G, s->planes, s->rho, mean_rho, s->P.inv_dx, s->P.inv_dy, s->d_scratch + 4, s->G.ywrap ? 0 : 1, s->st));
    s->launches++;
    double h[8];
    CUDA_TRY(cudaMemcpyAsync(h, s->d_scratch, 64, cudaMemcpyDeviceToHost, s->st));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    *out = h[4];
    return 0;
}

// tiles of this slab owned by `rank` (all when rank < 0), in gid order, with the attributes saveRankData writes
static void dump_tile_list(mpic_sim* s, int rank, std::vector<int>& gids, std::vector<int>& rows, std::vector<int>& cols, std::vector<int>& cur) {
    const int gtc = s->P.gtc;
    for (int r = s->prm.row0; r < s->prm.row0 + s->prm.nrows; r++)
        for (int c = 0; c < gtc; c++) {
            const int gid = r * gtc + c;
            if (rank >= 0 && s->owner[gid] != rank) continue;
            gids.push_back(gid); rows.push_back(r); cols.push_back(c); cur.push_back(s->owner[gid]);
        }
}
int mpic_dump_fields(mpic_sim* s, const char* filename, int rank) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && filename, MPIC_EINVAL, "bad argument");
    std::vector<int> gids, rows, cols, cur;
    dump_tile_list(s, rank, gids, rows, cols, cur);
    const size_t per = (size_t)(s->P.tnx + 4) * (s->P.tny + 4) * 6;
    std::vector<double> grid(per * gids.size());
    if (!gids.empty()) TRY(mpic_download_tiles(s, grid.data(), gids.data(), (int)gids.size()));
    return mpic_save_rank_data(filename, (int)gids.size(), gids.data(), rows.data(), cols.data(), cur.data(), s->P.tnx, s->P.tny,
                               s->prm.guard, grid.data());
}
}  // extern "C"

// ---- asynchronous dump: phase H of the loop (PIC_2D.cpp:414-419) off the step path.  The reference stops all ranks while
// saveRankData copies tile after tile into libhdf5; here the patches are formed by one kernel on the step stream (the only
// part the steps wait for: one read of the planes), travel to pinned host memory on a side stream, and a writer thread builds
// the file -- both overlap the following steps.
struct DumpState {
    cudaStream_t st = nullptr;
    cudaEvent_t ev_patches = nullptr, ev_host = nullptr;
    double* d_grid = nullptr; size_t d_bytes = 0;
    int* d_lists = nullptr; size_t l_bytes = 0;
    double* h_grid = nullptr; size_t h_bytes = 0;       // pinned
    std::vector<int> gids, rows, cols, cur, lists;
    std::string filename, err;
    std::thread worker;
    bool busy = false;
    int status = 0;
    int tnx = 0, tny = 0, guard = 0, device = 0;
};
static void dump_release(mpic_sim* s) {
    DumpState* d = s->dump;
    if (!d) return;
    if (d->worker.joinable()) d->worker.join();
    cudaFree(d->d_grid); cudaFree(d->d_lists); cudaFreeHost(d->h_grid);
    if (d->ev_patches) cudaEventDestroy(d->ev_patches);
    if (d->ev_host) cudaEventDestroy(d->ev_host);
    if (d->st) cudaStreamDestroy(d->st);
    delete d;
    s->dump = nullptr;
}
extern "C" {
int mpic_dump_wait(mpic_sim* s) {
    REQUIRE(s, MPIC_EINVAL, "null handle");
    DumpState* d = s->dump;
    if (!d || !d->busy) return 0;
    if (d->worker.joinable()) d->worker.join();
    d->busy = false;
    if (d->status) mpic_set_error("%s", d->err.c_str());
    return d->status;
}
int mpic_dump_fields_async(mpic_sim* s, const char* filename, int rank) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && filename, MPIC_EINVAL, "bad argument");
    if (!s->dump) {
        DumpState* d = new DumpState();
        s->dump = d;
        CUDA_TRY(cudaStreamCreateWithFlags(&d->st, cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreateWithFlags(&d->ev_patches, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&d->ev_host, cudaEventDisableTiming));
    }
    DumpState* d = s->dump;
    TRY(mpic_dump_wait(s));                              // one dump in flight: the staging buffers are single
    if (s->peer_on) TRY(mpic_sync(s));                   // guard rows are written by the neighbours: await their last phase
    d->gids.clear(); d->rows.clear(); d->cols.clear(); d->cur.clear();
    dump_tile_list(s, rank, d->gids, d->rows, d->cols, d->cur);
    const int nt = (int)d->gids.size();
    d->filename = filename; d->tnx = s->P.tnx; d->tny = s->P.tny; d->guard = s->prm.guard; d->device = s->prm.device;
    d->status = 0; d->err.clear();
    const size_t bytes = (size_t)nt * (s->P.tnx + 4) * (s->P.tny + 4) * 48;
    if (nt > 0) {
        d->lists.resize(2 * (size_t)nt);
        for (int t = 0; t < nt; t++) { d->lists[t] = d->rows[t] - s->prm.row0; d->lists[nt + t] = d->cols[t]; }
        TRY(mpic_scratch((void**)&d->d_lists, &d->l_bytes, d->lists.size() * 4));
        TRY(mpic_scratch((void**)&d->d_grid, &d->d_bytes, bytes));
        if (d->h_bytes < bytes) {
            if (d->h_grid) cudaFreeHost(d->h_grid);
            d->h_grid = nullptr; d->h_bytes = 0;
            CUDA_TRY(cudaMallocHost(&d->h_grid, bytes));
            d->h_bytes = bytes;
        }
        // the list outlives the copy (member of the dump state); the patch kernel is the only piece on the step stream
        CUDA_TRY(cudaMemcpyAsync(d->d_lists, d->lists.data(), d->lists.size() * 4, cudaMemcpyHostToDevice, s->st));
        TRY(launch_planes_to_tiles(s->G, s->planes, d->d_grid, d->d_lists, d->d_lists + nt, nt, s->P.tnx, s->P.tny, s->st));
        s->launches++;
        CUDA_TRY(cudaEventRecord(d->ev_patches, s->st));
        CUDA_TRY(cudaStreamWaitEvent(d->st, d->ev_patches, 0));
        CUDA_TRY(cudaMemcpyAsync(d->h_grid, d->d_grid, bytes, cudaMemcpyDeviceToHost, d->st));
    }
    CUDA_TRY(cudaEventRecord(d->ev_host, d->st));
    d->busy = true;
    d->worker = std::thread([d, nt]() {
        cudaSetDevice(d->device);
        cudaError_t e = cudaEventSynchronize(d->ev_host);
        if (e != cudaSuccess) { d->status = MPIC_ECUDA; d->err = std::string("async dump: ") + cudaGetErrorString(e); return; }
        const int rc = mpic_save_rank_data(d->filename.c_str(), nt, d->gids.data(), d->rows.data(), d->cols.data(), d->cur.data(),
                                           d->tnx, d->tny, d->guard, d->h_grid);
        if (rc) { d->status = rc; d->err = mpic_last_error(); }
    });
    return 0;
}

int mpic_selftest_math(int64_t n, uint64_t seed, int64_t* mismatches) {
    REQUIRE(n > 0 && mismatches, MPIC_EINVAL, "bad argument");
    TRY(need_device());
    unsigned long long* d;
    CUDA_TRY(cudaMalloc(&d, 8));
    CUDA_TRY(cudaMemset(d, 0, 8));
    int r = launch_selftest_math(n, seed, d, 0);
    unsigned long long h = 0;
    if (!r) { cudaError_t e = cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost); if (e != cudaSuccess) { mpic_set_error("selftest: %s", cudaGetErrorString(e)); r = MPIC_ECUDA; } }
    cudaFree(d);
    *mismatches = (int64_t)h;
    return r;
}

int mpic_get_owner(mpic_sim* s, int* owner) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && owner, MPIC_EINVAL, "bad argument");
    memcpy(owner, s->owner.data(), s->owner.size() * sizeof(int));
    return 0;
}
int mpic_set_owner(mpic_sim* s, const int* owner) {
    REQUIRE(s && owner, MPIC_EINVAL, "bad argument");
    memcpy(s->owner.data(), owner, s->owner.size() * sizeof(int));
    return 0;
}
int mpic_migrate_tile(mpic_sim* s, int gid, int dest) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && gid >= 0 && gid < (int)s->owner.size() && dest >= 0, MPIC_EINVAL, "bad tile/rank");
    s->owner[gid] = dest;     // on one device the fields and particles of the tile do not move
    return 0;
}

}  // extern "C"
