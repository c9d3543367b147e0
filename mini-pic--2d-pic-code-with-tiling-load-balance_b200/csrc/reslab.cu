// reslab.cu -- run-time load balancing on the device: cost-driven re-partition of the tile rows over the slabs.
//
// Replaces the reference's tile hand-off removeTileFromRank / addTileToRank (Auxiliar_functions.cpp:242-272: MPI_Send / MPI_Recv
// of ONE tile's grid; the particle vector is dropped) and its hard-coded trigger (PIC_2D.cpp:399-412: ranks 0-2
// give local tile #3 to rank 3 at step 5).  The policy -- which tile rows go where -- is the one the reference leaves as future
// work (Report p.10): minimise the most loaded slab under cost = particles + w_cell * cells.
//
// Why this is cheap here: particles are kept sorted by (tile, cell), so a tile row is ONE contiguous slot range of each of the
// seven particle arrays, and one contiguous block of rows of each field plane.  A rank packs what each new owner gets into a
// parcel (device memory), exports it (CUDA IPC handle; a plain pointer inside one process) and the new owner pulls it over
// NVLink with device-to-device copies: no host staging, no new handle -- the handle is rebuilt IN PLACE on its new rows.
// The parcels of a slab, concatenated in row order, are already in the sort order of the new slab; only the cell offsets are
// recomputed.  The current accumulators are zero between steps and the guard rows are rebuilt (x: wrap; y: the ring's refresh),
// so fields + particles are the whole state.  Results are partition-independent bit for bit (fixed-point current).
#include "sim.cuh"
#include <unistd.h>
#include <string.h>
#include <algorithm>

namespace {

constexpr unsigned long long PARCEL_MAGIC = 0x4d5049435041524bull;   // "MPICPARK"
struct ParcelBlob {                   // what travels inside mpic_parcel_info
    unsigned long long magic;
    int pid, device, row_lo, row_hi, nx, tny, has_ipc, pad1;   // has_ipc: h is valid (a process that cannot export still serves handles of its own process)
    unsigned long long nonce;         // mpic_process_nonce() of the packer: raw is trusted only when pid AND nonce are ours
    int64_t np;                       // particles in the parcel
    int64_t bytes;
    void* raw;                        // valid inside the packing process
    cudaIpcMemHandle_t h;
};
static_assert(sizeof(ParcelBlob) <= sizeof(mpic_parcel_info), "parcel blob does not fit");

// parcel layout: [6 planes][rows * tny][nx] doubles | x | y | ux | uy | uz | q (np doubles each) | tag (np int32)
inline int64_t parcel_field_values(const ParcelBlob& b) { return (int64_t)6 * (b.row_hi - b.row_lo) * b.tny * b.nx; }
inline int64_t parcel_bytes(const ParcelBlob& b) { return (parcel_field_values(b) + 6 * b.np) * 8 + ((b.np * 4 + 15) & ~15ll); }

}  // namespace

extern "C" {

int mpic_get_rows(mpic_sim* s, int* row0, int* nrows) {
    REQUIRE(s && row0 && nrows, MPIC_EINVAL, "bad argument");
    *row0 = s->prm.row0; *nrows = s->prm.nrows;
    return 0;
}

// ------------------------------------------------------------------------------------------ policy (host, exact)
int mpic_partition_rows(int n, const double* c, int nranks, int* rows_out) {
    REQUIRE(c && rows_out && n >= 1 && nranks >= 1, MPIC_EINVAL, "bad argument");
    REQUIRE(nranks <= n, MPIC_EINVAL, "more ranks (%d) than tile rows (%d)", nranks, n);
    // greedy sweep under a bottleneck limit; every slab still to be opened keeps one row
    auto cuts = [&](double limit, std::vector<int>& out) {
        out.clear();
        int start = 0;
        double acc = 0.0;
        for (int i = 0; i < n; i++) {
            const int remaining = nranks - (int)out.size() / 2 - 1;
            const bool must_cut = (n - i) == remaining && i > start;
            if (i > start && (acc + c[i] > limit || must_cut)) { out.push_back(start); out.push_back(i - start); start = i; acc = 0.0; }
            acc += c[i];
        }
        out.push_back(start); out.push_back(n - start);
    };
    double lo = c[0], hi = 0.0;
    for (int i = 0; i < n; i++) { REQUIRE(c[i] >= 0.0, MPIC_EINVAL, "negative row cost"); lo = c[i] > lo ? c[i] : lo; hi += c[i]; }
    std::vector<int> parts;
    for (int it = 0; it < 60; it++) {
        const double mid = 0.5 * (lo + hi);
        cuts(mid, parts);
        if ((int)parts.size() / 2 <= nranks) hi = mid; else lo = mid;
    }
    cuts(hi, parts);
    while ((int)parts.size() / 2 < nranks) {            // split the widest slab until every rank owns rows
        int k = 0;
        for (int j = 0; j < (int)parts.size() / 2; j++) if (parts[2 * j + 1] > parts[2 * k + 1]) k = j;
        const int r0 = parts[2 * k], nr = parts[2 * k + 1];
        REQUIRE(nr >= 2, MPIC_EINVAL, "cannot give every rank a tile row");
        parts[2 * k + 1] = nr / 2;
        parts.insert(parts.begin() + 2 * k + 2, { r0 + nr / 2, nr - nr / 2 });
    }
    for (int k = 0; k < 2 * nranks; k++) rows_out[k] = parts[k];
    return 0;
}

int mpic_rebalance_plan(int gtr, int nranks, const int64_t* row_particles, int64_t cells_per_tile_row, double w_cell,
                        double min_gain, const int* old_rows, int* new_rows, double cost[2]) {
    REQUIRE(row_particles && old_rows && new_rows && gtr >= 1 && nranks >= 1 && cells_per_tile_row > 0 && w_cell >= 0.0, MPIC_EINVAL, "bad argument");
    std::vector<double> c((size_t)gtr);
    for (int r = 0; r < gtr; r++) c[r] = (double)row_particles[r] + w_cell * (double)cells_per_tile_row;
    int covered = 0;
    for (int k = 0; k < nranks; k++) { REQUIRE(old_rows[2 * k] == covered && old_rows[2 * k + 1] >= 1, MPIC_EINVAL, "old_rows is not a partition"); covered += old_rows[2 * k + 1]; }
    REQUIRE(covered == gtr, MPIC_EINVAL, "old_rows does not cover the %d tile rows", gtr);
    TRY(mpic_partition_rows(gtr, c.data(), nranks, new_rows));
    auto worst = [&](const int* rows) {
        double w = 0.0;
        for (int k = 0; k < nranks; k++) { double a = 0.0; for (int r = rows[2 * k]; r < rows[2 * k] + rows[2 * k + 1]; r++) a += c[r]; w = a > w ? a : w; }
        return w;
    };
    const double w_old = worst(old_rows), w_new = worst(new_rows);
    if (cost) { cost[0] = w_old; cost[1] = w_new; }
    const bool same = memcmp(old_rows, new_rows, sizeof(int) * 2 * (size_t)nranks) == 0;
    if (same || w_new > (1.0 - min_gain) * w_old) { memcpy(new_rows, old_rows, sizeof(int) * 2 * (size_t)nranks); return 0; }
    return 1;
}

int mpic_get_tile_row_counts(mpic_sim* s, int64_t* counts) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && counts, MPIC_EINVAL, "bad argument");
    const int nt = s->P.ltr * s->P.gtc;
    if (s->cap == 0) { for (int r = 0; r < s->P.ltr; r++) counts[r] = 0; return 0; }
    std::vector<int> tc((size_t)nt);
    TRY(mpic_get_tile_counts(s, tc.data()));
    for (int r = 0; r < s->P.ltr; r++) { int64_t a = 0; for (int c = 0; c < s->P.gtc; c++) a += tc[(size_t)r * s->P.gtc + c]; counts[r] = a; }
    return 0;
}

// ------------------------------------------------------------------------------------------ mechanism
int mpic_reslab_pack(mpic_sim* s, int row_lo, int row_hi, mpic_parcel_info* out) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && out, MPIC_EINVAL, "bad argument");
    REQUIRE(row_lo >= s->prm.row0 && row_hi <= s->prm.row0 + s->prm.nrows && row_lo < row_hi, MPIC_EINVAL,
            "rows [%d,%d) are not on this slab [%d,%d)", row_lo, row_hi, s->prm.row0, s->prm.row0 + s->prm.nrows);
    TRY(mpic_ensure_sorted(s));
    REQUIRE(s->arrived == 0 && !s->counts_valid, MPIC_ESTATE, "re-partition between steps only (after the tile sort)");
    const FieldGeom& G = s->G;
    const PicConst& P = s->P;
    ParcelBlob b;
    memset(&b, 0, sizeof b);
    b.magic = PARCEL_MAGIC; b.pid = (int)getpid(); b.nonce = mpic_process_nonce(); b.device = s->prm.device; b.row_lo = row_lo; b.row_hi = row_hi; b.nx = G.nx; b.tny = P.tny;
    int64_t s_lo = 0;
    if (s->cap > 0 && s->n > 0) {
        if (s->perm_valid) {                             // physical tile/cell order (one gather pass), cell offsets stay valid
            TRY(launch_gather(s->pa[s->cur], s->pa[s->cur ^ 1], s->perm, s->n, nullptr, s->st));
            s->launches++;
            s->cur ^= 1; s->perm_valid = false;
        }
        int se[2];
        const int64_t c_lo = (int64_t)(row_lo - s->prm.row0) * P.gtc * P.cpt, c_hi = (int64_t)(row_hi - s->prm.row0) * P.gtc * P.cpt;
        CUDA_TRY(cudaMemcpyAsync(&se[0], s->cell_start + c_lo, 4, cudaMemcpyDeviceToHost, s->st));
        CUDA_TRY(cudaMemcpyAsync(&se[1], s->cell_start + c_hi, 4, cudaMemcpyDeviceToHost, s->st));
        CUDA_TRY(cudaStreamSynchronize(s->st));
        s_lo = se[0]; b.np = se[1] - se[0];
        REQUIRE(b.np >= 0 && se[1] <= s->n, MPIC_ESTATE, "cell offsets are stale (slots %d..%d of %lld)", se[0], se[1], (long long)s->n);
    }
    b.bytes = parcel_bytes(b);
    void* buf = nullptr;
    CUDA_TRY(cudaMalloc(&buf, (size_t)b.bytes));
    s->parcels.push_back(buf);
    double* f = (double*)buf;
    const int nyr = (row_hi - row_lo) * P.tny, j0 = (row_lo - s->prm.row0) * P.tny;
    for (int p = 0; p < 6; p++)
        CUDA_TRY(cudaMemcpy2DAsync(f + (int64_t)p * nyr * G.nx, (size_t)G.nx * 8, s->planes + p * G.plane + (int64_t)(G.g + j0) * G.pitch + MPIC_X0,
                                   (size_t)G.pitch * 8, (size_t)G.nx * 8, nyr, cudaMemcpyDeviceToDevice, s->st));
    if (b.np > 0) {
        const Particles& pa = s->pa[s->cur];
        const double* src[6] = { pa.x, pa.y, pa.ux, pa.uy, pa.uz, pa.q };
        double* d = f + parcel_field_values(b);
        for (int k = 0; k < 6; k++) CUDA_TRY(cudaMemcpyAsync(d + (int64_t)k * b.np, src[k] + s_lo, (size_t)b.np * 8, cudaMemcpyDeviceToDevice, s->st));
        CUDA_TRY(cudaMemcpyAsync(d + 6 * b.np, pa.tag + s_lo, (size_t)b.np * 4, cudaMemcpyDeviceToDevice, s->st));
    }
    CUDA_TRY(cudaStreamSynchronize(s->st));              // the parcel is complete before its handle leaves this call
    b.raw = buf;
    if (cudaIpcGetMemHandle(&b.h, buf) == cudaSuccess) b.has_ipc = 1;
    else cudaGetLastError();                              // single-process rings only need the raw pointer
    memset(out, 0, sizeof *out);
    memcpy(out->bytes, &b, sizeof b);
    return 0;
}

int mpic_reslab_apply(mpic_sim* s, int new_row0, int new_nrows, int64_t new_max_particles, const mpic_parcel_info* parcels,
                      int nparcels, const int* all_rows) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && parcels && nparcels >= 1, MPIC_EINVAL, "bad argument");
    REQUIRE(s->prm.nranks > 1, MPIC_ESTATE, "a single slab owns every tile row: nothing to re-partition");
    std::vector<ParcelBlob> bl((size_t)nparcels);
    for (int k = 0; k < nparcels; k++) {
        memcpy(&bl[k], parcels[k].bytes, sizeof(ParcelBlob));
        REQUIRE(bl[k].magic == PARCEL_MAGIC, MPIC_EINVAL, "parcel %d is not a parcel blob", k);
        REQUIRE(bl[k].nx == s->prm.nx && bl[k].tny == s->prm.tny, MPIC_EINVAL, "parcel %d comes from a different grid", k);
    }
    std::sort(bl.begin(), bl.end(), [](const ParcelBlob& a, const ParcelBlob& b) { return a.row_lo < b.row_lo; });
    int next = new_row0;
    int64_t total = 0;
    for (const ParcelBlob& b : bl) {
        REQUIRE(b.row_lo == next && b.row_hi > b.row_lo, MPIC_EINVAL, "parcels do not tile the rows [%d,%d): gap or overlap at row %d",
                new_row0, new_row0 + new_nrows, next);
        next = b.row_hi; total += b.np;
    }
    REQUIRE(next == new_row0 + new_nrows, MPIC_EINVAL, "parcels end at row %d, the slab at %d", next, new_row0 + new_nrows);
    int64_t cap = new_max_particles;
    if (cap == 0 && (s->prm.max_particles > 0 || total > 0)) cap = total <= s->prm.max_particles ? s->prm.max_particles : total + total / 4 + 4096;
    REQUIRE(total <= cap, MPIC_ENOMEM, "particle capacity %lld exceeded by the new population (%lld)", (long long)cap, (long long)total);
    mpic_params np_ = s->prm;
    np_.row0 = new_row0; np_.nrows = new_nrows; np_.max_particles = cap;
    {
        const int gtr = np_.ny / np_.tny;
        REQUIRE(new_row0 >= 0 && new_nrows >= 1 && new_row0 + new_nrows <= gtr, MPIC_EINVAL, "bad slab rows");
        REQUIRE((int64_t)np_.nx * new_nrows * np_.tny < (1ll << 31) - 8 && cap < (1ll << 31) - 64, MPIC_EINVAL, "slab too large for 32-bit indices");
    }
    // ---- the old state goes (nobody maps it any more: every rank disconnected before the blobs travelled) ...
    CUDA_TRY(cudaStreamSynchronize(s->st));
    if (s->dump) TRY(mpic_dump_wait(s));
    const int64_t launches = s->launches;
    mpic_free_state(s, cap);
    // ---- ... and the new one comes
    s->prm = np_;
    TRY(mpic_set_geometry(s));
    TRY(mpic_alloc_state(s));
    s->launches = launches;
    const FieldGeom& G = s->G;
    std::vector<void*> opened;
    int64_t at = 0;
    int rc = 0;
    for (const ParcelBlob& b : bl) {
        const void* src = b.raw;
        if (b.pid != (int)getpid() || b.nonce != mpic_process_nonce()) {
            if (!b.has_ipc) { mpic_set_error("parcel rows %d..%d come from another process that could not export a CUDA IPC handle", b.row_lo, b.row_hi); rc = MPIC_ECUDA; break; }
            void* ptr = nullptr;
            cudaError_t e = cudaIpcOpenMemHandle(&ptr, b.h, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) { mpic_set_error("cudaIpcOpenMemHandle(parcel rows %d..%d): %s", b.row_lo, b.row_hi, cudaGetErrorString(e)); rc = MPIC_ECUDA; break; }
            opened.push_back(ptr);
            src = ptr;
        } else if (b.device != s->prm.device) {          // same process, other GPU: direct peer copies when the link allows
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, s->prm.device, b.device) == cudaSuccess && can) cudaDeviceEnablePeerAccess(b.device, 0);
            cudaGetLastError();
        }
        const double* f = (const double*)src;
        const int nyr = (b.row_hi - b.row_lo) * b.tny, j0 = (b.row_lo - new_row0) * b.tny;
        cudaError_t e = cudaSuccess;
        for (int p = 0; p < 6 && e == cudaSuccess; p++)
            e = cudaMemcpy2DAsync(s->planes + p * G.plane + (int64_t)(G.g + j0) * G.pitch + MPIC_X0, (size_t)G.pitch * 8,
                                  f + (int64_t)p * nyr * G.nx, (size_t)G.nx * 8, (size_t)G.nx * 8, nyr, cudaMemcpyDefault, s->st);
        if (e == cudaSuccess && b.np > 0) {
            Particles& pa = s->pa[s->cur];
            double* dst[6] = { pa.x, pa.y, pa.ux, pa.uy, pa.uz, pa.q };
            const double* d = f + parcel_field_values(b);
            for (int k = 0; k < 6 && e == cudaSuccess; k++) e = cudaMemcpyAsync(dst[k] + at, d + (int64_t)k * b.np, (size_t)b.np * 8, cudaMemcpyDefault, s->st);
            if (e == cudaSuccess) e = cudaMemcpyAsync(pa.tag + at, d + 6 * b.np, (size_t)b.np * 4, cudaMemcpyDefault, s->st);
            at += b.np;
        }
        if (e != cudaSuccess) { mpic_set_error("parcel copy (rows %d..%d): %s", b.row_lo, b.row_hi, cudaGetErrorString(e)); rc = MPIC_ECUDA; break; }
    }
    if (!rc) rc = mpic_refresh_x_guards(s);
    if (!rc && s->cap > 0) { s->n = total; rc = mpic_sort_uploaded(s); }
    if (!rc && cudaStreamSynchronize(s->st) != cudaSuccess) { mpic_set_error("re-partition: %s", cudaGetErrorString(cudaGetLastError())); rc = MPIC_ECUDA; }
    for (void* p : opened) cudaIpcCloseMemHandle(p);
    TRY(rc);
    if (all_rows) {
        const int gtc = s->P.gtc;
        for (int k = 0; k < s->prm.nranks; k++)
            for (int r = all_rows[2 * k]; r < all_rows[2 * k] + all_rows[2 * k + 1]; r++)
                for (int c = 0; c < gtc; c++) if ((size_t)r * gtc + c < s->owner.size()) s->owner[(size_t)r * gtc + c] = k;
    }
    return 0;
}

int mpic_reslab_finish(mpic_sim* s) {
    if (!s) return 0;
    cudaSetDevice(s->prm.device);
    for (void* p : s->parcels) cudaFree(p);
    s->parcels.clear();
    return 0;
}

}  // extern "C"
