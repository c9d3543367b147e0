// peer.cu -- the slab ring over peer-mapped device memory (NVLink / NVSwitch): no staging, no copies, no host barriers.
//
// Replaces, on the step path of a multi-GPU run, the reference's exchange phases (PIC_2D.cpp:200-248,276-319,348-396:
// packSendBuffer -> MPI_Isend/Irecv -> updateGuardRegion, two MPI_Barriers each) and the particle hand-off that the
// reference's tile migration lacks (Auxiliar_functions.cpp:244-246 drops particles).  Each handle maps its two ring
// neighbours' field planes, current accumulators, arrival buffers and control block (CUDA IPC between the torchrun
// processes; plain pointers when the handles share a process) and then
//   * the Yee sweeps (yee.cu, store_images) write their edge rows straight into the neighbours' guard rows,
//   * peer_jsend_kernel adds the current guard rows into the neighbours' edge rows with remote 64-bit integer atomics
//     (order-independent, so still bit-identical to the single-slab result),
//   * peer_ship_kernel writes the leaver records and their count into the neighbours' arrival buffers,
//   * peer_sync_kernel orders the phases: release-store of an epoch into both neighbours' control blocks, acquire-spin on
//     the own control block (bounded by a timeout that raises an error word instead of hanging the GPU).
// Why no write-after-read hazard exists on the guard rows: every remote write of phase P is issued only after the
// neighbour signalled the phase that last READ those rows (J before the first half-B, B1 before E, E before the second
// half-B, B2 before the next push) -- see DESIGN.md section 4.
#include "sim.cuh"
#include <unistd.h>
#include <string.h>

namespace {

enum Phase { PH_J = 0, PH_B1 = 1, PH_E = 2, PH_B2 = 3, PH_R1 = 4, PH_R2 = 5, PH_F = 6, PH_COUNT = 7 };   // PH_F: fused field step
constexpr int CTL_WORDS = 64;
constexpr int CTL_ARRIVE = 16;        // + parity*2 + from_side : record count written by that neighbour
constexpr int CTL_ERROR = 24;         // non-zero: a wait timed out (value = 1 + phase)
constexpr int CTL_DONE = 32;          // two 32-bit edge-CTA counters of the fused sweep ordering (local use)
constexpr unsigned long long PEER_MAGIC = 0x4d50494350454552ull;   // "MPICPEER"

struct PeerBlob {                     // what travels inside mpic_peer_info
    unsigned long long magic;
    int pid, device, rank, nyl, lcap, pad;
    unsigned long long nonce;         // mpic_process_nonce() of the exporter: raw[] is trusted only when pid AND nonce are ours
    int64_t plane;
    void* raw[5];                     // planes, jacc, arrive block, control block, planes_alt (valid inside the exporting process)
    cudaIpcMemHandle_t h[5];
    unsigned long long has[5];
};
static_assert(sizeof(PeerBlob) <= sizeof(mpic_peer_info), "peer blob does not fit");

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long gtimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// thread 0 <-> lower neighbour, thread 1 <-> upper neighbour: signal `sig` (all remote stores of the kernels before this
// one on the stream are ordered before it), then wait until that neighbour signalled `want`.  After the J phase the same
// threads hand the migration counts to the host through mapped pinned memory (host_out: leave[2], arrive[2], error) --
// no device-to-host copy sits on the stream.
__global__ void peer_sync_kernel(unsigned long long* lo_flag, unsigned long long* hi_flag, unsigned long long sig,
                                 const unsigned long long* my_flags, unsigned long long want, unsigned long long* err,
                                 int phase, unsigned long long timeout_ns, const int* leave_cnt,
                                 const unsigned long long* arrive_cnt, volatile long long* host_out) {
    const int t = threadIdx.x;
    if (t > 1) return;
    unsigned long long* out = t ? hi_flag : lo_flag;
    __threadfence_system();
    if (sig) st_release_sys(out, sig);
    if (want) {
        const unsigned long long t0 = gtimer_ns();
        while (ld_acquire_sys(my_flags + t) < want) {
            if (gtimer_ns() - t0 > timeout_ns) { atomicExch(err, 1ull + phase); break; }
            __nanosleep(32);
        }
    }
    if (host_out) {
        host_out[t] = leave_cnt[t];
        host_out[2 + t] = (long long)ld_acquire_sys(arrive_cnt + t);
        __syncwarp(0x3);
        if (t == 0) host_out[4] = (long long)ld_acquire_sys(err);
        __threadfence_system();
    }
}

// After the push, ONE launch hands everything that crossed the slab edges to the neighbours:
//   blockIdx.y <  6g : current guard rows -> neighbours' interior edge rows (remote 64-bit ADD), own guard rows zeroed
//                      (blockIdx.y = side*3g + plane*g + r)
//   blockIdx.y >= 6g : leaver records (written by the push kernel) -> the neighbours' arrival buffers, plus the count
//                      (blockIdx.y - 6g = direction)
__device__ __forceinline__ void handoff_done(unsigned int* done, unsigned long long* sig_lo, unsigned long long* sig_hi,
                                             unsigned long long sig) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        if (atomicAdd(done, 1u) == gridDim.x * gridDim.y - 1) {       // every CTA's remote stores are ordered before the flags
            *done = 0;
            st_release_sys(sig_lo, sig);
            st_release_sys(sig_hi, sig);
        }
    }
}
__global__ void peer_handoff_kernel(FieldGeom G, unsigned long long* __restrict__ jacc, unsigned long long* lo_j, int64_t lo_plane,
                                    int lo_nyl, unsigned long long* hi_j, int64_t hi_plane, LeaveBufs lv, double* lo_dst,
                                    unsigned long long* lo_cnt, int lo_cap, double* hi_dst, unsigned long long* hi_cnt, int hi_cap,
                                    unsigned int* done, unsigned long long* sig_lo, unsigned long long* sig_hi, unsigned long long sig) {
    const int g = G.g;
    if ((int)blockIdx.y < 6 * g) {
        const int r = blockIdx.y % g, p = (blockIdx.y / g) % 3, side = blockIdx.y / (3 * g);
        for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < G.pitch; c += gridDim.x * blockDim.x) {
            unsigned long long* mine = jacc + (int64_t)p * G.plane + (int64_t)(side ? g + G.nyl + r : r) * G.pitch + c;
            const unsigned long long v = *mine;
            if (!v) continue;
            unsigned long long* dst = side ? hi_j + (int64_t)p * hi_plane + (int64_t)(g + r) * G.pitch + c
                                           : lo_j + (int64_t)p * lo_plane + (int64_t)(lo_nyl + r) * G.pitch + c;
            atomicAdd(dst, v);
            *mine = 0ull;
        }
        handoff_done(done, sig_lo, sig_hi, sig);
        return;
    }
    const int dir = blockIdx.y - 6 * g;
    const int raw = lv.count[dir];
    const int cap = dir ? hi_cap : lo_cap;
    const int n = raw < lv.cap ? (raw < cap ? raw : cap) : (lv.cap < cap ? lv.cap : cap);
    double* dst = dir ? hi_dst : lo_dst;
    const double* src = lv.rec[dir];
    const int64_t total = (int64_t)n * MPIC_REC_DOUBLES;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) dst[e] = src[e];
    if (blockIdx.x == 0 && threadIdx.x == 0) *(dir ? hi_cnt : lo_cnt) = (unsigned long long)raw;
    handoff_done(done, sig_lo, sig_hi, sig);
}

// migration counts -> host (mapped pinned memory); launched behind the first half-B sweep, whose edge CTAs awaited both J flags
__global__ void peer_counts_kernel(const int* leave_cnt, const unsigned long long* arrive_cnt, const unsigned long long* err,
                                   volatile long long* host_out) {
    const int t = threadIdx.x;
    if (t < 2) { host_out[t] = leave_cnt[t]; host_out[2 + t] = (long long)ld_acquire_sys(arrive_cnt + t); }
    if (t == 2) host_out[4] = (long long)ld_acquire_sys(err);
    __threadfence_system();
}

// whole edge rows (x guards included) of `nplanes` planes -> neighbours' guard rows; used after uploads
__global__ void peer_halo_push_kernel(FieldGeom G, const double* __restrict__ f, int nplanes, PeerHalo H) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= G.pitch) return;
    const int g = G.g, r = blockIdx.y % g, p = (blockIdx.y / g) % nplanes, side = blockIdx.y / (nplanes * g);
    const double v = f[(int64_t)p * G.plane + (int64_t)(side ? G.nyl + r : g + r) * G.pitch + c];
    if (side) H.hi[(int64_t)p * H.hi_plane + (int64_t)r * G.pitch + c] = v;
    else H.lo[(int64_t)p * H.lo_plane + (int64_t)(H.lo_row0 + r) * G.pitch + c] = v;
}

}  // namespace

#define LAUNCH_CHECK() do { cudaError_t _e = cudaGetLastError(); if (_e != cudaSuccess) { mpic_set_error("%s:%d: launch failed: %s", __FILE__, __LINE__, cudaGetErrorString(_e)); return MPIC_ECUDA; } } while (0)

namespace {

PeerHalo halo_of(const mpic_sim* s) {
    PeerHalo h;
    h.lo = s->peer[0].planes; h.hi = s->peer[1].planes;
    h.lo_plane = s->peer[0].plane; h.hi_plane = s->peer[1].plane;
    h.lo_row0 = s->G.g + s->peer[0].nyl;
    return h;
}

unsigned long long timeout_ns() {
    static unsigned long long t = 0;
    if (!t) { const char* e = getenv("MPIC_PEER_TIMEOUT_MS"); t = (unsigned long long)(e ? atoll(e) : 8000) * 1000000ull; }
    return t;
}

// signal phase `ph` with the next epoch to both neighbours and/or wait for theirs.  Signal and wait may be issued apart
// (B2: signalled after the sweep, awaited only before the next push, so the tile sort in between absorbs any skew).
int phase_launch(mpic_sim* s, int ph, bool signal, bool wait, bool counts = false, int par = 0, cudaStream_t on = nullptr) {
    if (signal) ++s->epoch[ph];
    const unsigned long long e = s->epoch[ph];
    // I am the lower neighbour's UPPER side (word 1) and the upper neighbour's LOWER side (word 0)
    peer_sync_kernel<<<1, 32, 0, on ? on : s->st>>>(s->peer[0].ctl + ph * 2 + 1, s->peer[1].ctl + ph * 2 + 0, signal ? e : 0ull, s->ctl + ph * 2,
                                          wait ? e : 0ull, s->ctl + CTL_ERROR, ph, timeout_ns(), counts ? s->lv.count : nullptr,
                                          s->ctl + CTL_ARRIVE + par * 2, counts ? (volatile long long*)s->h_counts : nullptr);
    LAUNCH_CHECK();
    s->launches++;
    return 0;
}
int phase_sync(mpic_sim* s, int ph) { return phase_launch(s, ph, true, true); }
int b2_wait_if_pending(mpic_sim* s) {
    if (!s->b2_pending) return 0;
    s->b2_pending = false;
    return phase_launch(s, s->pending_ph, false, true);
}

const char* k_trace_names[12] = { "b2wait", "push", "handoff", "syncJ", "fold_x", "B1", "syncB1", "E", "syncE", "B2+signal",
                                  "host+absorb+sort", "" };
inline void tp(mpic_sim* s, int k) { if (s->trace) cudaEventRecord(s->tev[k], s->st); }
void trace_collect(mpic_sim* s) {                     // after the step's last event has been recorded
    if (!s->trace) return;
    cudaEventSynchronize(s->tev[11]);
    for (int k = 0; k < 11; k++) { float ms = 0; if (cudaEventElapsedTime(&ms, s->tev[k], s->tev[k + 1]) == cudaSuccess) s->tacc[k] += ms; }
    s->tsteps++;
}
void trace_report(mpic_sim* s) {
    if (!s->trace || !s->tsteps) return;
    fprintf(stderr, "[mpic peer trace] rank %d, %lld steps, mean us per phase:", s->prm.rank, s->tsteps);
    double tot = 0;
    for (int k = 0; k < 11; k++) { fprintf(stderr, " %s=%.1f", k_trace_names[k], 1e3 * s->tacc[k] / s->tsteps); tot += s->tacc[k]; }
    fprintf(stderr, " total=%.1f\n", 1e3 * tot / s->tsteps);
}

int check_error_word(mpic_sim* s, unsigned long long w) {
    REQUIRE(w == 0, MPIC_ESTATE, "peer ring: rank %d timed out waiting for a neighbour in phase %d (a rank died, or the ranks "
            "do not call the same sequence)", s->prm.rank, (int)w - 1);
    return 0;
}

int open_side(mpic_sim* s, int side, const PeerBlob& b, const mpic_sim::PeerSide* same) {
    mpic_sim::PeerSide& p = s->peer[side];
    p.plane = b.plane; p.nyl = b.nyl; p.lcap = b.lcap;
    if (same) { p.planes = same->planes; p.planes_alt = same->planes_alt; p.jacc = same->jacc; p.arrive = same->arrive; p.ctl = same->ctl; p.ipc = false; return 0; }
    void* ptr[5] = { nullptr, nullptr, nullptr, nullptr, nullptr };
    if (b.pid == (int)getpid() && b.nonce == mpic_process_nonce()) {   // same process: plain pointers (+ peer access across devices)
        if (b.device != s->prm.device) {
            int can = 0;
            CUDA_TRY(cudaDeviceCanAccessPeer(&can, s->prm.device, b.device));
            REQUIRE(can, MPIC_ESTATE, "device %d cannot access device %d", s->prm.device, b.device);
            cudaError_t e = cudaDeviceEnablePeerAccess(b.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CUDA_TRY(e);
            cudaGetLastError();
        }
        for (int k = 0; k < 5; k++) ptr[k] = b.raw[k];
    } else {
        for (int k = 0; k < 5; k++)
            if (b.has[k]) CUDA_TRY(cudaIpcOpenMemHandle(&ptr[k], b.h[k], cudaIpcMemLazyEnablePeerAccess));
        p.ipc = true;
    }
    p.planes = (double*)ptr[0]; p.planes_alt = (double*)ptr[4]; p.jacc = (unsigned long long*)ptr[1]; p.arrive = (double*)ptr[2]; p.ctl = (unsigned long long*)ptr[3];
    return 0;
}

}  // namespace

// unmap the neighbours; the handle's own peer-visible memory (control block, planes, ...) stays: the neighbours may still hold
// mappings of it until THEY disconnect
static void peer_unmap(mpic_sim* s) {
    for (int side = 0; side < 2; side++) {
        mpic_sim::PeerSide& p = s->peer[side];
        if (p.ipc) { cudaIpcCloseMemHandle(p.planes); if (p.planes_alt) cudaIpcCloseMemHandle(p.planes_alt); cudaIpcCloseMemHandle(p.jacc); if (p.arrive) cudaIpcCloseMemHandle(p.arrive); cudaIpcCloseMemHandle(p.ctl); }
        p = mpic_sim::PeerSide();
    }
    s->peer_on = false;
    s->b2_pending = false;
}
void mpic_peer_release(mpic_sim* s) {
    if (!s) return;
    peer_unmap(s);
    trace_report(s);
    for (int k = 0; k < 16; k++) if (s->tev[k]) { cudaEventDestroy(s->tev[k]); s->tev[k] = nullptr; }
    s->trace = false;
    if (s->h_counts) { cudaFreeHost(s->h_counts); s->h_counts = nullptr; }
    if (s->ev_counts) { cudaEventDestroy(s->ev_counts); s->ev_counts = nullptr; }
    if (s->ev_hand) { cudaEventDestroy(s->ev_hand); s->ev_hand = nullptr; }
    if (s->ctl) { cudaFree(s->ctl); s->ctl = nullptr; }
}

extern "C" {

int mpic_peer_export(mpic_sim* s, mpic_peer_info* out) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && out, MPIC_EINVAL, "bad argument");
    REQUIRE(s->prm.nranks > 1, MPIC_ESTATE, "the peer ring exists only for nranks > 1");
    if (!s->ctl) {
        CUDA_TRY(cudaMalloc(&s->ctl, CTL_WORDS * 8));
        CUDA_TRY(cudaMemset(s->ctl, 0, CTL_WORDS * 8));
        CUDA_TRY(cudaMallocHost(&s->h_counts, 16 * sizeof(long long)));
        CUDA_TRY(cudaEventCreateWithFlags(&s->ev_counts, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&s->ev_hand, cudaEventDisableTiming));
    }
    memset(out, 0, sizeof *out);
    PeerBlob b;
    memset(&b, 0, sizeof b);
    b.magic = PEER_MAGIC; b.pid = (int)getpid(); b.nonce = mpic_process_nonce(); b.device = s->prm.device; b.rank = s->prm.rank; b.nyl = s->G.nyl;
    b.lcap = s->lv.cap; b.plane = s->G.plane;
    // The second plane block of the fused field step is peer-visible too -- when there is one.  It is only of use to scheme 0
    // (mpic_peer_field_step_fused refuses the others), it costs another six planes of HBM (3.2 GB on an 8192^2 slab), and a ring
    // that never takes a fused step does not need it: MPIC_PEER_NO_FUSED=1 leaves it out.  If it cannot be set up (out of memory,
    // tensor-map encoding) the plain ring must still work: the block is then simply not exported, connect copes with has[4] == 0
    // and mpic_peer_field_step_fused says "neighbours exported no second plane block".
    if (s->prm.scheme == 0 && !getenv("MPIC_PEER_NO_FUSED") && mpic_fused_setup(s) != 0) {
        cudaGetLastError();
        if (s->planes_alt && !s->fused_ready) { cudaFree(s->planes_alt); s->planes_alt = nullptr; }
    }
    b.raw[0] = s->planes; b.raw[1] = s->jacc; b.raw[2] = s->arrive_blk; b.raw[3] = s->ctl; b.raw[4] = s->fused_ready ? s->planes_alt : nullptr;
    for (int k = 0; k < 5; k++)
        if (b.raw[k]) { CUDA_TRY(cudaIpcGetMemHandle(&b.h[k], b.raw[k])); b.has[k] = 1; }
    memcpy(out->bytes, &b, sizeof b);
    return 0;
}

int mpic_peer_disconnect(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s, MPIC_EINVAL, "null handle");
    if (!s->peer_on) return 0;
    // everything this handle still has in flight towards its neighbours (edge rows, flags) lands before the mappings go
    CUDA_TRY(cudaStreamSynchronize(s->st));
    unsigned long long w = 0;
    CUDA_TRY(cudaMemcpy(&w, s->ctl + CTL_ERROR, 8, cudaMemcpyDeviceToHost));
    peer_unmap(s);
    return check_error_word(s, w);
}

int mpic_peer_connect(mpic_sim* s, const mpic_peer_info* lo, const mpic_peer_info* hi) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && lo && hi, MPIC_EINVAL, "bad argument");
    REQUIRE(s->ctl, MPIC_ESTATE, "call mpic_peer_export first");
    PeerBlob bl, bh;
    memcpy(&bl, lo->bytes, sizeof bl); memcpy(&bh, hi->bytes, sizeof bh);
    REQUIRE(bl.magic == PEER_MAGIC && bh.magic == PEER_MAGIC, MPIC_EINVAL, "not a peer blob");
    const int n = s->prm.nranks;
    REQUIRE(bl.rank == (s->prm.rank + n - 1) % n && bh.rank == (s->prm.rank + 1) % n, MPIC_EINVAL,
            "rank %d: neighbours must be ranks %d and %d (got %d, %d)", s->prm.rank, (s->prm.rank + n - 1) % n, (s->prm.rank + 1) % n, bl.rank, bh.rank);
    TRY(open_side(s, 0, bl, nullptr));
    TRY(open_side(s, 1, bh, bh.rank == bl.rank ? &s->peer[0] : nullptr));       // two ranks: both neighbours are the same handle
    // no kernel may be loaded lazily once flag waits can spin (see yee_preload)
    TRY(yee_preload());
    TRY(particles_preload());
    {
        cudaFuncAttributes a;
        const void* fs[] = { (const void*)peer_sync_kernel, (const void*)peer_handoff_kernel, (const void*)peer_halo_push_kernel,
                             (const void*)peer_counts_kernel };
        for (const void* f : fs) CUDA_TRY(cudaFuncGetAttributes(&a, f));
    }
    if (const char* e = getenv("MPIC_PEER_TRACE")) if (atoi(e) > 0 && !s->trace) {
        for (int k = 0; k < 12; k++) CUDA_TRY(cudaEventCreate(&s->tev[k]));
        s->trace = true;
    }
    s->peer_on = true;
    return 0;
}

// per-phase device times of the peer PIC step (CUDA events on the handle's stream; costs one extra host wait per step, so
// it is switched on only around a measurement).  enable: 1 = start / reset, 0 = stop, -1 = leave as is.  out_ms[11]: mean ms
// per step of {b2wait, push, handoff, -, -, B1+counts, -, E, -, B2, host+absorb+sort}; *steps = steps accumulated.
int mpic_peer_trace(mpic_sim* s, int enable, double* out_ms, long long* steps) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && s->peer_on, MPIC_ESTATE, "peer ring not connected");
    if (out_ms) for (int k = 0; k < 11; k++) out_ms[k] = s->tsteps ? s->tacc[k] / (double)s->tsteps : 0.0;
    if (steps) *steps = s->tsteps;
    if (enable == 1) {
        for (int k = 0; k < 12; k++) if (!s->tev[k]) CUDA_TRY(cudaEventCreate(&s->tev[k]));
        for (int k = 0; k < 16; k++) s->tacc[k] = 0.0;
        s->tsteps = 0;
        s->trace = true;
    } else if (enable == 0) s->trace = false;
    return 0;
}

int mpic_peer_refresh_begin(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && s->peer_on, MPIC_ESTATE, "peer ring not connected");
    TRY(b2_wait_if_pending(s));
    TRY(phase_sync(s, PH_R1));                           // both neighbours finished writing their own planes
    peer_halo_push_kernel<<<dim3(ceil_div(s->G.pitch, 256), 2 * 6 * s->G.g), 256, 0, s->st>>>(s->G, s->planes, 6, halo_of(s));
    LAUNCH_CHECK();
    s->launches++;
    return phase_sync(s, PH_R2);
}
int mpic_peer_refresh_end(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && s->peer_on, MPIC_ESTATE, "peer ring not connected");
    TRY(b2_wait_if_pending(s));
    unsigned long long w = 0;
    CUDA_TRY(cudaMemcpyAsync(&w, s->ctl + CTL_ERROR, 8, cudaMemcpyDeviceToHost, s->st));
    CUDA_TRY(cudaStreamSynchronize(s->st));
    return check_error_word(s, w);
}

// Phases A..F with the ordering fused into the sweeps: every sweep's edge CTAs await the flag of the phase that produced the
// guard rows they read (and that last read the guard rows they overwrite at the neighbour), the last edge CTA raises the
// sweep's own flag.  B1 awaits E of the previous step (a no-op after a B2, needed when scheme == 1 has no second half-B).
static PeerSync sweep_sync(mpic_sim* s, int wait_ph, int sig_ph) {
    PeerSync ps;
    ps.wait_flags = s->ctl + wait_ph * 2; ps.want = s->epoch[wait_ph];
    ps.sig = ++s->epoch[sig_ph];
    ps.sig_lo = s->peer[0].ctl + sig_ph * 2 + 1; ps.sig_hi = s->peer[1].ctl + sig_ph * 2 + 0;
    ps.done = (unsigned int*)(s->ctl + CTL_DONE);
    ps.err = s->ctl + CTL_ERROR; ps.timeout_ns = timeout_ns(); ps.phase = wait_ph;
    return ps;
}
static int yee_phases(mpic_sim* s, int with_j, int par = -1) {
    const PeerHalo h = halo_of(s);
    // in a PIC step the first half-B awaits the neighbours' J phase (they pushed, shipped, added their current -- and hence
    // finished reading the guard rows this sweep overwrites); in a field-only step, their last E sweep
    TRY(launch_b_sweep(s->G, s->planes, s->cf, h, sweep_sync(s, par >= 0 ? PH_J : PH_E, PH_B1), s->st));
    // PIC step: the E sweep reads the current -- complete once this rank's own hand-off (side stream) has zeroed / shipped the
    // guard rows and the neighbours' adds have landed (their J flags: acquired by the edge CTAs of the half-B just queued)
    if (par >= 0 && s->ev_hand && s->st_sort) CUDA_TRY(cudaStreamWaitEvent(s->st, s->ev_hand, 0));
    tp(s, 6); tp(s, 7);
    TRY(launch_e_sweep(s->G, s->planes, s->jacc, with_j, s->cf, h, sweep_sync(s, PH_B1, PH_E), s->st));
    tp(s, 8); tp(s, 9);
    s->launches += 2;
    if (s->prm.scheme == 0) {
        TRY(launch_b_sweep(s->G, s->planes, s->cf, h, sweep_sync(s, PH_E, PH_B2), s->st));
        s->launches++;
    }
    s->b2_pending = true;                                // the last sweep's flag is awaited before the next reader of the
    s->pending_ph = s->prm.scheme == 0 ? PH_B2 : PH_E;   // guard rows it fills at this rank (push, dump, download)
    tp(s, 10);
    return 0;
}

int mpic_peer_field_step(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && s->peer_on, MPIC_ESTATE, "peer ring not connected");
    TRY(b2_wait_if_pending(s));
    return yee_phases(s, 0);
}

// One fused kernel per field step over the ring: the slab-edge CTAs await the neighbours' PREVIOUS fused step (its stores
// filled the guard rows this step reads, and it was the last reader of the guard rows this step overwrites in the
// neighbours' other plane block), every other CTA starts at once; the last edge CTA of each side raises this step's flag.
// One flag exchange per step instead of three.  Every rank must take the same sequence of fused / unfused steps (the plane
// blocks swap in lockstep).
int mpic_peer_field_step_fused(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && s->peer_on, MPIC_ESTATE, "peer ring not connected");
    REQUIRE(s->prm.scheme == 0, MPIC_ESTATE, "fused step: scheme 0 only");
    REQUIRE(s->peer[0].planes_alt && s->peer[1].planes_alt, MPIC_ESTATE, "neighbours exported no second plane block");
    TRY(b2_wait_if_pending(s));                          // a preceding unfused phase: await it with its own flag
    PeerHalo h;                                          // images go into the neighbours' OTHER block (their dst of this step)
    h.lo = s->peer[0].planes_alt; h.hi = s->peer[1].planes_alt;
    h.lo_plane = s->peer[0].plane; h.hi_plane = s->peer[1].plane;
    h.lo_row0 = s->G.g + s->peer[0].nyl;
    TRY(launch_fused_step(s->G, &s->ftmap, s->planes_alt, s->jacc, 0, s->cf, h, sweep_sync(s, PH_F, PH_F), s->st));
    s->launches++;
    mpic_fused_swap(s);
    s->b2_pending = true;                                // the flag just raised by the neighbours is awaited by the next reader
    s->pending_ph = PH_F;
    return 0;
}

int mpic_peer_step_begin(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && s->peer_on, MPIC_ESTATE, "peer ring not connected");
    REQUIRE(s->cap > 0 && s->arrive_blk, MPIC_ESTATE, "no particle storage");
    const int par = (int)(++s->step_no & 1);
    const FieldGeom& G = s->G;
    tp(s, 0);
    if (s->b2_pending) {                                 // awaited by the push kernel's slab-edge CTAs themselves
        s->b2_pending = false;
        s->lv.wait_flags = s->ctl + s->pending_ph * 2; s->lv.wait_want = s->epoch[s->pending_ph];
        s->lv.wait_err = s->ctl + CTL_ERROR; s->lv.wait_timeout_ns = timeout_ns(); s->lv.wait_phase = s->pending_ph;
    }
    tp(s, 1);
    // Main stream: push -> half-B -> E -> half-B.  Side stream (the sort stream), behind the push: leaver records -> hand-off
    // (J guard rows added into the neighbours, leavers shipped, J flag raised) -> wait for the neighbours' J flags and hand the
    // migration counts to the host -> (mpic_peer_step_end) histogram, absorb, stable sort.  Only the E sweep waits for the
    // hand-off; everything that concerns particles runs next to the three field sweeps.
    cudaStream_t side = s->st_sort ? s->st_sort : s->st;
    s->defer_count = s->st_sort != nullptr;              // push: no histogram / leaver launches on the main stream
    const int rc = mpic_push_deposit(s);
    s->defer_count = false;
    s->lv.wait_want = 0;
    TRY(rc);
    tp(s, 2);
    if (side != s->st) {
        CUDA_TRY(cudaStreamWaitEvent(side, s->ev_push, 0));
        TRY(launch_leavers(s->P, s->pa[s->cur], s->cell_start, s->dirs, s->mcnt, s->lv_off, s->cell_count, s->lv, side));
        s->launches += 2;
    }
    const int64_t rec = MPIC_REC_DOUBLES;
    // ONE launch hands over the current guard rows (remote adds) and the leaver records, and its last CTA raises the J flag.
    // My leavers going DOWN arrive at the lower neighbour "from above" (side 1); going UP: at the upper neighbour's side 0.
    const int gx = ceil_div(G.pitch, 256) > 16 ? ceil_div(G.pitch, 256) : 16;
    const unsigned long long ej = ++s->epoch[PH_J];
    peer_handoff_kernel<<<dim3(gx, 6 * G.g + 2), 256, 0, side>>>(G, (unsigned long long*)s->jacc, s->peer[0].jacc, s->peer[0].plane,
        s->peer[0].nyl, s->peer[1].jacc, s->peer[1].plane, s->lv,
        s->peer[0].arrive + (int64_t)(par * 2 + 1) * s->peer[0].lcap * rec, s->peer[0].ctl + CTL_ARRIVE + par * 2 + 1, s->peer[0].lcap,
        s->peer[1].arrive + (int64_t)(par * 2 + 0) * s->peer[1].lcap * rec, s->peer[1].ctl + CTL_ARRIVE + par * 2 + 0, s->peer[1].lcap,
        (unsigned int*)(s->ctl + CTL_DONE) + 2, s->peer[0].ctl + PH_J * 2 + 1, s->peer[1].ctl + PH_J * 2 + 0, ej);
    LAUNCH_CHECK();
    s->launches++;
    if (side != s->st) CUDA_TRY(cudaEventRecord(s->ev_hand, side));
    // both neighbours' hand-offs (their J flags) -> leaver / arrival counts to the host; the one host wait of the step is taken
    // on this event in mpic_peer_step_end
    TRY(phase_launch(s, PH_J, false, true, true, par, side));
    CUDA_TRY(cudaEventRecord(s->ev_counts, side));
    tp(s, 3); tp(s, 4); tp(s, 5);
    return yee_phases(s, 1, par);
}

int mpic_peer_step_end(mpic_sim* s) {
    if (s) cudaSetDevice(s->prm.device);
    REQUIRE(s && s->peer_on, MPIC_ESTATE, "peer ring not connected");
    REQUIRE(s->counts_valid, MPIC_ESTATE, "mpic_peer_step_end must follow mpic_peer_step_begin");
    CUDA_TRY(cudaEventSynchronize(s->ev_counts));
    const int par = (int)(s->step_no & 1);
    const long long* hc = (const long long*)s->h_counts;
    const unsigned long long arr[2] = { (unsigned long long)hc[2], (unsigned long long)hc[3] }, err = (unsigned long long)hc[4];
    TRY(check_error_word(s, err));
    const int leave[2] = { (int)hc[0], (int)hc[1] };
    for (int d = 0; d < 2; d++)
        REQUIRE(leave[d] <= s->lv.cap && leave[d] <= s->peer[d].lcap && (int64_t)arr[d] <= s->lv.cap, MPIC_ENOMEM,
                "migration buffer overflow: %d leavers / %lld arrivals > capacity %d", leave[d], (long long)arr[d], s->lv.cap);
    s->left = (int64_t)leave[0] + leave[1];
    s->migrated_total += s->left;
    REQUIRE(s->n + (int64_t)arr[0] + (int64_t)arr[1] <= s->cap, MPIC_ENOMEM, "particle capacity exceeded by arrivals");
    // The counts are known once the first half-B is done; the E sweep and the second half-B are still running on the main stream.
    // Histogram, absorb and the stable sort touch particle arrays and cell tables only, the sweeps planes and current only: they
    // go to the sort stream, ordered behind the counts kernel, and the main stream waits for them at the end.
    cudaStream_t ss = s->st_sort ? s->st_sort : s->st;
    if (ss != s->st) CUDA_TRY(cudaStreamWaitEvent(ss, s->ev_counts, 0));
    if (!s->moves_counted) {
        TRY(launch_count_moves(s->P, s->G.ywrap, s->mcnt, s->cell_count, ss));
        s->launches++;
        s->moves_counted = true;
    }
    for (int side = 0; side < 2; side++) {               // appended in (lower, upper) order -- SURVEY.md B.8
        const int n = (int)arr[side];
        TRY(launch_absorb(s->P, s->pa[s->cur], s->n + s->arrived, s->arrive_blk + (int64_t)(par * 2 + side) * s->lv.cap * MPIC_REC_DOUBLES,
                          n, s->cell_count, s->keys, s->G.err, ss));
        s->arrived += n;
        s->launches += n > 0;
    }
    TRY(mpic_sort_after_push_on(s, ss));
    tp(s, 11);
    trace_collect(s);
    return 0;
}

}  // extern "C"
