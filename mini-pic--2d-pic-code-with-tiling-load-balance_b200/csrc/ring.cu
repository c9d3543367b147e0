// ring.cu -- one host process driving N slab handles (one per GPU) as a periodic ring with CUDA peer copies over NVLink.
//
// This is the C++ driver's multi-GPU path (the torchrun/NCCL path with one process per GPU lives in slab.py; both use
// the same pack / unpack / leaver-record kernels and give bit-identical results).  It replaces the reference's exchange
// phases (PIC_2D.cpp:200-248,276-319,348-396) and tile hand-off (Auxiliar_functions.cpp:242-272) for a slab
// decomposition: per exchange every handle packs its two edge messages, each neighbour pulls its message with
// cudaMemcpyPeerAsync on ITS OWN stream after waiting on the sender's "packed" event, and unpacks.  No host
// synchronisation and no barrier on the field path; the sender's next pack waits on the receivers' "copied" events
// (write-after-read on the staging buffer).  Handles may share a device (used by the one-GPU tests).
#include "sim.cuh"
#include <algorithm>

namespace {

struct RingEvents {
    cudaEvent_t packed = nullptr, copied = nullptr;
};
// events live beside the handle; created lazily (one pair per handle)
RingEvents* events_of(mpic_sim* s) {
    if (!s->ring_events) {
        RingEvents* e = new RingEvents();
        cudaSetDevice(s->prm.device);
        cudaEventCreateWithFlags(&e->packed, cudaEventDisableTiming);
        cudaEventCreateWithFlags(&e->copied, cudaEventDisableTiming);
        cudaEventRecord(e->copied, s->st);
        s->ring_events = e;
    }
    return (RingEvents*)s->ring_events;
}

int check_ring(mpic_sim* const* sims, int n) {
    REQUIRE(sims && n >= 2, MPIC_EINVAL, "ring needs >= 2 handles");
    for (int r = 0; r < n; r++) {
        REQUIRE(sims[r] && sims[r]->prm.nranks == n && sims[r]->prm.rank == r, MPIC_EINVAL,
                "ring handle %d must be rank %d of %d", r, r, n);
        REQUIRE(sims[r]->halo_send, MPIC_ESTATE, "handle %d has no halo staging", r);
    }
    return 0;
}

// move `bytes` from src (device pointer on handle a) to dst (on handle b) on b's stream, after a's `packed` event
int pull(mpic_sim* b, void* dst, mpic_sim* a, const void* src, size_t bytes) {
    if (!bytes) return 0;
    CUDA_TRY(cudaSetDevice(b->prm.device));
    CUDA_TRY(cudaMemcpyPeerAsync(dst, b->prm.device, src, a->prm.device, bytes, b->st));
    return 0;
}

}  // namespace

void mpic_ring_release(mpic_sim* s) {
    if (s && s->ring_events) {
        RingEvents* e = (RingEvents*)s->ring_events;
        cudaEventDestroy(e->packed); cudaEventDestroy(e->copied);
        delete e;
        s->ring_events = nullptr;
    }
}

extern "C" {

int mpic_ring_enable_peer_access(mpic_sim* const* sims, int n) {
    TRY(check_ring(sims, n));
    for (int r = 0; r < n; r++)
        for (int d = -1; d <= 1; d += 2) {
            const int a = sims[r]->prm.device, b = sims[(r + d + n) % n]->prm.device;
            if (a == b) continue;
            int can = 0;
            CUDA_TRY(cudaDeviceCanAccessPeer(&can, a, b));
            if (!can) continue;                         // cudaMemcpyPeerAsync still works (staged by the driver)
            CUDA_TRY(cudaSetDevice(a));
            cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CUDA_TRY(e);
            cudaGetLastError();
        }
    return 0;
}

int mpic_ring_exchange(mpic_sim* const* sims, int n, int which) {
    TRY(check_ring(sims, n));
    REQUIRE(which >= 0 && which <= 2, MPIC_EINVAL, "bad which");
    // 1. pack (after the neighbours finished reading the previous message out of the staging buffer)
    for (int r = 0; r < n; r++) {
        mpic_sim* s = sims[r];
        CUDA_TRY(cudaSetDevice(s->prm.device));
        for (int d = -1; d <= 1; d += 2) CUDA_TRY(cudaStreamWaitEvent(s->st, events_of(sims[(r + d + n) % n])->copied, 0));
        TRY(mpic_halo_pack(s, which));
        CUDA_TRY(cudaEventRecord(events_of(s)->packed, s->st));
    }
    // 2. every handle pulls: from the lower neighbour its send[1] (towards upper) -> my recv[0]; from the upper its send[0] -> my recv[1]
    const size_t bytes = (size_t)sims[0]->halo_side * 8;
    for (int r = 0; r < n; r++) {
        mpic_sim* s = sims[r];
        mpic_sim* lo = sims[(r - 1 + n) % n];
        mpic_sim* hi = sims[(r + 1) % n];
        CUDA_TRY(cudaSetDevice(s->prm.device));
        CUDA_TRY(cudaStreamWaitEvent(s->st, events_of(lo)->packed, 0));
        CUDA_TRY(cudaStreamWaitEvent(s->st, events_of(hi)->packed, 0));
        TRY(pull(s, s->halo_recv, lo, lo->halo_send + lo->halo_side, bytes));
        TRY(pull(s, s->halo_recv + s->halo_side, hi, hi->halo_send, bytes));
        CUDA_TRY(cudaEventRecord(events_of(s)->copied, s->st));
        TRY(mpic_halo_unpack(s, which));
    }
    return 0;
}

int mpic_ring_refresh_guards(mpic_sim* const* sims, int n) {
    TRY(mpic_ring_exchange(sims, n, 0));
    return mpic_ring_exchange(sims, n, 1);
}

int mpic_ring_field_steps(mpic_sim* const* sims, int n, int nsteps) {
    TRY(check_ring(sims, n));
    for (int k = 0; k < nsteps; k++) {
        for (int r = 0; r < n; r++) TRY(mpic_step_b_half(sims[r]));
        TRY(mpic_ring_exchange(sims, n, 1));
        for (int r = 0; r < n; r++) TRY(mpic_step_e_full(sims[r]));
        TRY(mpic_ring_exchange(sims, n, 0));
        if (sims[0]->prm.scheme == 0) {
            for (int r = 0; r < n; r++) TRY(mpic_step_b_half(sims[r]));
            TRY(mpic_ring_exchange(sims, n, 1));
        }
    }
    return 0;
}

// particle hand-off: one host sync per handle to learn the counts, then peer copies of the records
int mpic_ring_migrate(mpic_sim* const* sims, int n) {
    TRY(check_ring(sims, n));
    std::vector<int> leave(2 * (size_t)n), arrive(2 * (size_t)n);
    for (int r = 0; r < n; r++) TRY(mpic_migration_counts(sims[r], &leave[2 * r], &arrive[2 * r]));
    for (int r = 0; r < n; r++) {
        mpic_sim* s = sims[r];
        const int lo = (r - 1 + n) % n, hi = (r + 1) % n;
        REQUIRE(arrive[2 * r] == leave[2 * lo + 1] && arrive[2 * r + 1] == leave[2 * hi], MPIC_ESTATE,
                "leaver counts in the halo header disagree with the neighbours' (rank %d)", r);
        // the senders are idle (synchronised above), so no event is needed before pulling
        TRY(pull(s, s->arrive[0], sims[lo], sims[lo]->lv.rec[1], (size_t)arrive[2 * r] * MPIC_REC_DOUBLES * 8));
        TRY(pull(s, s->arrive[1], sims[hi], sims[hi]->lv.rec[0], (size_t)arrive[2 * r + 1] * MPIC_REC_DOUBLES * 8));
        TRY(mpic_migration_absorb(s, arrive[2 * r], arrive[2 * r + 1]));
        CUDA_TRY(cudaEventRecord(events_of(s)->copied, s->st));
    }
    // the next push (which rewrites the senders' record buffers) must wait for the pulls
    for (int r = 0; r < n; r++) {
        mpic_sim* s = sims[r];
        CUDA_TRY(cudaSetDevice(s->prm.device));
        for (int d = -1; d <= 1; d += 2) CUDA_TRY(cudaStreamWaitEvent(s->st, events_of(sims[(r + d + n) % n])->copied, 0));
    }
    return 0;
}

int mpic_ring_pic_steps(mpic_sim* const* sims, int n, int nsteps) {
    TRY(check_ring(sims, n));
    for (int k = 0; k < nsteps; k++) {
        for (int r = 0; r < n; r++) TRY(mpic_push_deposit(sims[r]));
        TRY(mpic_ring_exchange(sims, n, 2));
        TRY(mpic_ring_field_steps(sims, n, 1));
        TRY(mpic_ring_migrate(sims, n));
        for (int r = 0; r < n; r++) TRY(mpic_sort_tiles(sims[r]));
    }
    return 0;
}

// Cost-driven re-partition of the ring (reslab.cu holds policy and mechanism; this is the sequence for handles that share a
// process): occupancy per tile row -> plan -> every handle packs one parcel per new owner -> every handle is rebuilt in place
// from the parcels of its new rows (device-to-device / peer copies) -> parcels freed -> y guards refreshed over the ring.
int mpic_ring_rebalance(mpic_sim* const* sims, int n, double w_cell, double min_gain, int* changed, int* rows_out) {
    TRY(check_ring(sims, n));
    const mpic_params& p0 = sims[0]->prm;
    const int gtr = p0.ny / p0.tny;
    std::vector<int64_t> rowp((size_t)gtr, 0);
    std::vector<int> old_rows(2 * (size_t)n), new_rows(2 * (size_t)n);
    for (int r = 0; r < n; r++) {
        old_rows[2 * r] = sims[r]->prm.row0; old_rows[2 * r + 1] = sims[r]->prm.nrows;
        TRY(mpic_get_tile_row_counts(sims[r], rowp.data() + sims[r]->prm.row0));
    }
    double cost[2];
    const int plan = mpic_rebalance_plan(gtr, n, rowp.data(), (int64_t)p0.nx * p0.tny, w_cell, min_gain, old_rows.data(), new_rows.data(), cost);
    if (plan < 0) return plan;
    if (changed) *changed = plan;
    if (rows_out) for (int k = 0; k < 2 * n; k++) rows_out[k] = new_rows[k];
    if (!plan) return 0;
    // owner[] follows the slabs only where it described them before (the default); a caller-defined map of logical ranks
    // (host/pic2d --ranks R: dump grouping) is left alone
    bool owner_is_slab = true;
    for (int r = 0; r < n && owner_is_slab; r++) {
        const int gtc = sims[r]->P.gtc;
        for (int k = 0; k < n && owner_is_slab; k++)
            for (int row = old_rows[2 * k]; row < old_rows[2 * k] + old_rows[2 * k + 1] && owner_is_slab; row++)
                for (int c = 0; c < gtc; c++) if (sims[r]->owner[(size_t)row * gtc + c] != k) { owner_is_slab = false; break; }
    }
    for (int r = 0; r < n; r++) TRY(mpic_peer_disconnect(sims[r]));     // a peer-memory ring over the same handles: unmap first
    std::vector<std::vector<mpic_parcel_info>> inbox((size_t)n);
    for (int r = 0; r < n; r++)
        for (int d = 0; d < n; d++) {
            const int lo = std::max(old_rows[2 * r], new_rows[2 * d]);
            const int hi = std::min(old_rows[2 * r] + old_rows[2 * r + 1], new_rows[2 * d] + new_rows[2 * d + 1]);
            if (hi <= lo) continue;
            mpic_parcel_info info;
            TRY(mpic_reslab_pack(sims[r], lo, hi, &info));
            inbox[d].push_back(info);
        }
    for (int r = 0; r < n; r++)
        TRY(mpic_reslab_apply(sims[r], new_rows[2 * r], new_rows[2 * r + 1], 0, inbox[r].data(), (int)inbox[r].size(), owner_is_slab ? new_rows.data() : nullptr));
    for (int r = 0; r < n; r++) TRY(mpic_reslab_finish(sims[r]));
    TRY(mpic_ring_enable_peer_access(sims, n));
    return mpic_ring_refresh_guards(sims, n);
}

}  // extern "C"
