/* minipic_b200.h -- C ABI of the B200-native Mini-PIC time-step path (libminipic_b200.so).
 *
 * This is the drop-in boundary for the reference's per-timestep loop (PIC_2D.cpp:171-420).  The reference has no FFI: its
 * interface is three C++ headers statically linked into main().  Every entry point below names the
 * reference interface it replaces; host/ (C++) re-exports the reference's exact C++ signatures on top of this ABI and
 * INTEGRATION.md shows the binding a maintainer adds.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success or a negative MPIC_E* code (the reference
 * functions are all `void` and never check anything); no exceptions cross the boundary; one handle per GPU, one host thread
 * per handle.  All reals are IEEE binary64, all indices 32-bit int unless stated, exactly as in the reference.
 * There is NO CPU fallback: without a CUDA device every compute entry point returns MPIC_ENODEV.
 *
 * Two levels (SURVEY.md section 8b "Granularity"):
 *   L0  compat   one tile per call on HOST buffers, same semantics as Field_update.h / Auxiliar_functions.h.
 *   L1  resident device-resident simulation state advanced for ALL local tiles per call.
 */
#ifndef MINIPIC_B200_H
#define MINIPIC_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define MPIC_OK 0
#define MPIC_EINVAL (-1)   /* bad argument */
#define MPIC_ENODEV (-2)   /* no CUDA device / driver */
#define MPIC_ECUDA (-3)    /* CUDA runtime error (mpic_last_error() has the text) */
#define MPIC_ENOMEM (-4)   /* particle capacity exceeded */
#define MPIC_ESTATE (-5)   /* call sequence error (e.g. particles not loaded) */

const char* mpic_last_error(void);
int mpic_device_count(void);
/* ABI version; bumped whenever a signature or struct below changes. */
int mpic_abi_version(void);   /* currently 3 */

/* ============================================================ integer helpers (host, exact)
 * Replace Auxiliar_functions.h:55-65 (findBestGrid, findBestTileGrid, tileGlobalRow/Col, getGlobalID, getGlobalRowCol,
 * getNeighborGID, computeTag) and the tables dRow/dCol/opposite (Auxiliar_functions.h:11-13). */
void mpic_find_best_grid(int size, int* R, int* C);
void mpic_find_best_tile_grid(int num_tiles, int* tile_rows, int* tile_cols);
int mpic_tile_global_row(int rank_row, int tile_rows, int t_row);
int mpic_tile_global_col(int rank_col, int tile_cols, int t_col);
int mpic_get_global_id(int global_row, int global_col, int global_tile_cols);
void mpic_get_global_row_col(int gid, int global_tile_cols, int* row, int* col);
int mpic_get_neighbor_gid(int gid, int d, int global_tile_rows, int global_tile_cols);
int mpic_compute_tag(int gid, int d);
void mpic_direction_tables(int drow[8], int dcol[8], int opposite[8]);
/* slab decomposition over tile rows (north_star; R=N, C=1 instead of findBestGrid): rows [*row0, *row0 + *nrows) */
void mpic_slab_rows(int global_tile_rows, int nranks, int rank, int* row0, int* nrows);

/* ============================================================ L0 compat: one tile, host buffers
 * E3/B3: AoS triplets {Ex,Ey,Ez}/{Bx,By,Bz} (GridE/GridB, Field_update.h:9-17), tile incl. guards, row-major [j*TX+i].
 * grid6: AoS sextets (Grid, Auxiliar_functions.h:24-27).  Each call does H2D, one kernel, D2H. */
/* updateBhalf, Field_update.h:36 / Field_update.cpp:14-36: writes the interior of Bnew3 only. */
int mpic_update_b_half(const double* E3, const double* Bold3, double* Bnew3,
                       int interior_nx, int interior_ny, int guard, double dt, double dx, double dy);
/* updateEfull, Field_update.h:54 / Field_update.cpp:40-59: updates the interior of E3 in place. */
int mpic_update_e_full(double* E3, const double* B3,
                       int interior_nx, int interior_ny, int guard, double dt, double dx, double dy);
/* packSendBuffer, Auxiliar_functions.h:68 / .cpp:73-169: out6 gets mpic_strip_cells(d,...) cells; returns that count. */
int mpic_pack_send_buffer(const double* grid6, int interior_nx, int interior_ny, int guard, int d, double* out6);
/* updateGuardRegion, Auxiliar_functions.h:69 / .cpp:172-239. */
int mpic_update_guard_region(double* grid6, int interior_nx, int interior_ny, int guard, int d, const double* in6);
int mpic_strip_cells(int d, int interior_nx, int interior_ny, int guard);

/* ============================================================ L1 resident simulation */
typedef struct mpic_sim mpic_sim;

typedef struct mpic_params {
    int nx, ny;            /* GLOBAL interior cells (PIC_2D.cpp:62) */
    int tnx, tny;          /* tile interior cells (interior_nx/ny, PIC_2D.cpp:67-68); must divide nx, ny */
    int guard;             /* guard cells (PIC_2D.cpp:65); must be 2 */
    int scheme;            /* 0: current Field_update.cpp (half-B, E, half-B); 1: old_files v1 (B full dt, E) */
    int fma_mode;          /* 0: strict IEEE expression order (== reference built without contraction);
                              1: per-statement FMA contraction (== the build that produced the shipped dumps) */
    int rank, nranks;      /* slab index / count along y (whole tile rows) */
    int row0, nrows;       /* this slab's tile rows [row0,row0+nrows); nranks==1 -> 0, ny/tny */
    int device;            /* CUDA device ordinal */
    double dx, dy, dt;
    double qm;             /* species charge/mass (electrons: -1) */
    int64_t max_particles; /* capacity of each particle array (0: fields only); < 2^31 - 64 (32-bit slots; int32 tags) */
    int deposit;           /* 0: charge-conserving zigzag for Jx, Jy + CIC Jz (default; Gauss's law holds to round-off);
                              1: direct CIC of q v at the time-centred position on the E_k lattices (SURVEY.md B.5 option a, the
                                 literal reading of north_star; NOT charge conserving -- a selectable mode, not a tuned kernel) */
    int reserved;          /* 0 */
} mpic_params;
/* Fixed-point current: J is accumulated as int64 (value * 2^52), which makes the deposit order-independent and hence
 * bit-reproducible on any number of GPUs.  Range: |J| < 1024 code units per node and step; a step that comes within a factor
 * two of it raises an error that the next mpic_sync / mpic_upload_* returns as MPIC_EINVAL. */

int mpic_create(const mpic_params* p, mpic_sim** out);
int mpic_destroy(mpic_sim* s);
/* All work of a handle is issued on one stream; this blocks until it has drained. */
int mpic_sync(mpic_sim* s);
/* cudaStream_t of the handle (for event timing by the host harness). */
void* mpic_stream(mpic_sim* s);

/* ---- fields.  "planes": six (nyl, nx) row-major interior planes Ex,Ey,Ez,Bx,By,Bz of this slab, host memory.
 * Guards are rebuilt on the device (periodic in x; periodic in y when nranks==1, else left to mpic_halo_*). */
int mpic_upload_planes(mpic_sim* s, const double* const planes[6]);
int mpic_download_planes(mpic_sim* s, double* const planes[6]);
/* Reference tile layout: for each listed gid a (TY,TX) AoS Grid patch INCLUDING guards -- the layout of Tile::grid,
 * of the MPI wire format and of the HDF5 `fields` dataset (Auxiliar_functions.h:24-42, HDF5_output.cpp:30-46).
 * upload takes interiors only (guards are rebuilt); download fills guards from the periodic neighbours, i.e. what the
 * reference holds after an exchange phase (PIC_2D.cpp:200-248). */
int mpic_upload_tiles(mpic_sim* s, const double* grid6, const int* gids, int ntiles);
int mpic_download_tiles(mpic_sim* s, double* grid6, const int* gids, int ntiles);

/* Phase A/E (PIC_2D.cpp:174-196,322-344) for all local tiles + guard refresh (phase B/F). */
int mpic_step_b_half(mpic_sim* s);
/* Phase C (PIC_2D.cpp:251-272) incl. the -dt*J term when particles are present, + guard refresh (phase D). */
int mpic_step_e_full(mpic_sim* s);
/* nsteps field-only iterations of the loop PIC_2D.cpp:171-396 (A..F, or A..D for scheme 1). */
int mpic_field_steps(mpic_sim* s, int nsteps);
/* Same result (bit-identical) from ONE kernel per step (csrc/yee.cu fused_yee_kernel: tile + ring of all six planes staged
 * by one TMA box load, half-B -> E -> half-B in shared memory, out of place between two plane blocks): 96 B/cell of HBM
 * traffic instead of 216.  scheme 1 (two sweeps) falls back to mpic_field_steps. */
int mpic_field_steps_fused(mpic_sim* s, int nsteps);
/* Phases A..F of ONE step fused as above, with the -dt*J term (guard ring of J folded on the fly, accumulators zeroed)
 * when the handle holds particles.  Single slab, scheme 0.  mpic_pic_steps uses it. */
int mpic_step_fields_fused(mpic_sim* s);

/* ---- slab halos (nranks > 1): replace the inter-rank part of the exchange phases B/D/F (PIC_2D.cpp:200-248,276-319,
 * 348-396; MPI_Isend/Irecv of packSendBuffer strips) for a 1-D slab decomposition over tile rows.
 * which==0: E planes, 1: B planes, 2: current accumulators (int64; ADDED into the neighbour).
 * The library packs the edge rows into contiguous device staging buffers; the host moves send[d] to the neighbour's
 * recv[1-d] with any transport (NCCL send/recv through torch.distributed in slab.py, peer copies, MPI), then unpacks.
 * Side 0 = lower neighbour (rank-1, periodic), side 1 = upper neighbour (rank+1).  One staging set is shared by all three
 * `which`: finish one exchange (pack, transfer, unpack) before packing the next.  The which==2 message also carries,
 * in its 128-byte header, the number of particles that left towards that side in the last push. */
typedef struct mpic_halo_desc {
    void* send[2]; void* recv[2];   /* device pointers, each `bytes` long */
    int64_t bytes;                  /* bytes per message */
} mpic_halo_desc;
int mpic_halo_describe(mpic_sim* s, mpic_halo_desc* out);
int mpic_halo_pack(mpic_sim* s, int which);
/* fields: received rows -> guard rows.  current: received rows added into the interior edge rows, own guard rows
 * zeroed, then the x guards are folded (i.e. completes mpic_fold_current for a slab). */
int mpic_halo_unpack(mpic_sim* s, int which);

/* ---- particle migration between slabs (no reference counterpart; the reference's tile migration drops particles,
 * Auxiliar_functions.cpp:244-246).  mpic_push_deposit appends every particle that left the slab to send[0] (down) or
 * send[1] (up) as a 56-byte record {x,y,ux,uy,uz,q,tag}.  Protocol per step: push_deposit; halo exchange which==2;
 * field phases; mpic_migration_counts (one stream sync: own leaver counts + the neighbours', which arrived in the
 * which==2 header); transfer leave[d] records send[d] -> neighbour's recv[1-d]; mpic_migration_absorb; mpic_sort_tiles. */
typedef struct mpic_migr_desc {
    void* send[2]; void* recv[2];   /* device pointers to record arrays */
    int64_t cap;                    /* records per buffer */
    int64_t record_bytes;           /* 56 */
} mpic_migr_desc;
int mpic_migration_describe(mpic_sim* s, mpic_migr_desc* out);
int mpic_migration_counts(mpic_sim* s, int leave[2], int arrive[2]);
int mpic_migration_absorb(mpic_sim* s, int n_from_lo, int n_from_hi);
int64_t mpic_migrated_total(mpic_sim* s);   /* harness: records this slab has handed to its neighbours so far */

/* ---- peer-memory slab ring (nranks > 1; replaces, on the step path, the pack / transfer / unpack protocol above and
 * with it the reference's exchange phases PIC_2D.cpp:200-248,276-319,348-396 and MPI_Barriers :200,226,...).
 * Every handle maps its two ring neighbours' device memory (CUDA IPC between processes -- one process per GPU under
 * torchrun --, plain pointers inside one process) and the kernels talk over NVLink directly:
 *   - the Yee sweeps store their edge rows straight into the neighbours' guard rows (no pack, no copy, no unpack);
 *   - the current guard rows are added into the neighbours' edge rows with remote 64-bit integer atomics;
 *   - leaver records and their counts are written into the neighbours' arrival buffers;
 *   - phases are ordered by epoch flags in peer memory (release store / acquire spin), not by host barriers.
 * mpic_peer_info is an opaque, fixed-size, trivially copyable blob: move it between ranks with any transport
 * (torch.distributed all_gather in slab.py).  One host synchronisation per step remains (migration counts), placed so
 * that it overlaps the field sweeps. */
typedef struct mpic_peer_info { unsigned char bytes[512]; } mpic_peer_info;
int mpic_peer_export(mpic_sim* s, mpic_peer_info* out);
/* lo = rank-1, hi = rank+1 (periodic).  Blobs exported by handles of this very process are recognised (no IPC). */
int mpic_peer_connect(mpic_sim* s, const mpic_peer_info* lo, const mpic_peer_info* hi);
/* All ranks call these collectively (each on its own handle).  Split in two so that one host thread can drive several
 * handles: call begin on every handle, then end on every handle. */
int mpic_peer_refresh_begin(mpic_sim* s);   /* after uploads: y guards of E and B from the neighbours */
int mpic_peer_refresh_end(mpic_sim* s);
int mpic_peer_field_step(mpic_sim* s);      /* phases A..F of one step; no host synchronisation at all */
/* The same step (bit-identical) from the fused kernel of mpic_field_steps_fused: ONE launch and ONE flag exchange per
 * step; edge rows go straight into the neighbours' other plane block.  All ranks must call the same sequence. */
int mpic_peer_field_step_fused(mpic_sim* s);
int mpic_peer_step_begin(mpic_sim* s);      /* push+deposit, hand-off, J add, Yee phases: everything enqueued, no host wait */
int mpic_peer_step_end(mpic_sim* s);        /* migration counts (the one host wait), arrivals absorbed, tile sort */

/* Unmap both neighbours (closes the CUDA-IPC mappings, drops the flags); the handle keeps its own state.  Called on EVERY rank
 * before any rank frees or re-partitions memory the others have mapped (mpic_reslab_apply, mpic_destroy); idempotent. */
int mpic_peer_disconnect(mpic_sim* s);

/* Harness: per-phase device times of mpic_peer_step_begin/_end (events on the handle's stream; adds a host wait per step).
 * enable 1 = start/reset, 0 = stop, -1 = read only.  out_ms[11] = mean ms of {wait for the neighbours' last phase, push+deposit,
 * hand-off, -, -, first half-B, -, E, -, second half-B, migration counts + absorb + tile sort}. */
int mpic_peer_trace(mpic_sim* s, int enable, double* out_ms, long long* steps);

/* ---- particles (struct Particle, Auxiliar_functions.h:16-21, as SoA: x,y,px,py,pz,q doubles and int tag).
 * Upload accepts any order; the device sorts by (tile gid, cell) before the first push. */
int mpic_upload_particles(mpic_sim* s, int64_t n, const double* x, const double* y, const double* ux, const double* uy,
                          const double* uz, const double* q, const int* tag);
int mpic_download_particles(mpic_sim* s, int64_t cap, double* x, double* y, double* ux, double* uy, double* uz,
                            double* q, int* tag, int64_t* n_out);
int64_t mpic_num_particles(mpic_sim* s);
/* Device loader (plasma.py:load_uniform semantics, same bits). band_lo/hi in [0,1] fractions of box_y (0,1 = all). */
int mpic_load_uniform(mpic_sim* s, int ppc, double u_th, double drift_x, double drift_y, double drift_z,
                      double two_stream, double density, double band_lo, double band_hi, uint64_t seed, int append);

/* P1+P2: gather + Boris + move + charge-conserving deposit into the fixed-point J accumulators. */
int mpic_push_deposit(mpic_sim* s);
/* P3: fold the guard ring of the J accumulators into the interior (x always; y when nranks==1). */
int mpic_fold_current(mpic_sim* s);
/* S: per-cell count, prefix scan, STABLE sort into (tile gid, cell) order: inside a cell particles keep their
 * previous relative order, arrivals from the neighbour slabs follow in (lower, upper) neighbour order, each in the sender's
 * slot order -- no atomic decides a slot, so the order (what mpic_download_particles returns) is reproducible run to run and
 * equals a stable sort by cell key.  After a push only a permutation is built (9 B/particle); the next push gathers through it. */
int mpic_sort_tiles(mpic_sim* s);
/* nsteps full PIC iterations: push_deposit, J guard fold, A..F field phases with -dt*J, sort. */
int mpic_pic_steps(mpic_sim* s, int nsteps);

/* tile occupancy (local tiles, gid order within the slab) and per-particle tile gid in current slot order */
int mpic_get_tile_counts(mpic_sim* s, int* counts);
int mpic_get_particle_tile_index(mpic_sim* s, int* gid_of_particle);
/* Diagnostic: number of particles that changed CELL in the last push (slab leavers included).  Valid between
 * mpic_push_deposit and mpic_sort_tiles.  bench.py reports it as the regime's cell-crossing fraction. */
int mpic_cell_crossers(mpic_sim* s, int64_t* n_out);
/* J accumulators of the last push as int64 fixed point (value * 2^52), three (nyl,nx) planes */
int mpic_download_current(mpic_sim* s, int64_t* const jplanes[3]);
/* rho of the stored particles as int64 fixed point (value * 2^52; node weights of the Ez lattice), one (nyl,nx)
 * plane; guard contributions folded (x always; y on a single slab -- a slab's first row misses the lower neighbour's part). */
int mpic_download_rho(mpic_sim* s, int64_t* rho);
/* out[0] field energy 1/2 sum dx dy (E^2+B^2) over the slab; out[1] kinetic energy sum m_p (gamma-1);
 * out[2] max |div E - rho + <rho>| (Gauss residual); out[3] total charge; out[4] number of particles */
int mpic_diagnostics(mpic_sim* s, double out[5]);
/* max |div E - rho + mean_rho| over this slab with the caller's (GLOBAL) mean charge density: the multi-slab form of out[2]
 * (a slab's first cell row is skipped: its node charge needs the lower neighbour's particles). */
int mpic_gauss_residual(mpic_sim* s, double mean_rho, double* out);

/* ---- for callers that must not allocate device memory.  (No counterpart in the reference: it has no particles, no
 * diagnostics and no GPU.)
 * mpic_kinetic_energy_static computes what mpic_diagnostics reports as out[1] and out[3] -- out[0] = sum |q_p / qm| (gamma_p - 1),
 * the kinetic energy, and out[1] = sum q_p, the total charge, with the same reproducible ordered block sum and therefore the same
 * digits for the same particle order -- but WITHOUT a handle and without a single cudaMalloc / cudaFree, in this call or any
 * earlier one.  It is for the two kinds of caller that cannot afford mpic_create + mpic_diagnostics: a code whose particle
 * arrays live in its own device arena (own pool, torch's caching allocator, all of HBM reserved at start-up for a 1e9-particle
 * slab) with nothing left for the planes and sort tables a handle allocates; and a diagnostics hook between the steps of a
 * running peer ring, where cudaMalloc's device-wide synchronisation would stall every rank spinning on this rank's flags.
 *   ux, uy, uz, q   DEVICE pointers (memory of the device that is current in the calling thread), n doubles each;
 *   out             HOST array of two doubles;  qm: charge/mass ratio of the species (non-zero; electrons -1);
 *   stream          cudaStream_t to run on (NULL = the default stream); the call returns after that stream has drained.
 * The reduction's work area (block partials, ticket, result slots) is a static device array inside the library image,
 * MPIC_KINETIC_STATIC_SCRATCH_BYTES long, resident on every GPU that loads the library until the process exits.  It does not
 * grow with n; the one size limit is n <= MPIC_KINETIC_STATIC_MAX_PARTICLES (the library's 32-bit particle-slot range), beyond
 * which the call returns MPIC_EINVAL and says so -- reduce a larger population in pieces and add the results.
 * The first call on each device looks up the array's device address (cudaGetSymbolAddress; this may load the library's device
 * code): make that first call at set-up time, before any stream capture is begun on that device.
 * ALL calls on a device share the one work area, and it keeps the previous call's contents (the call clears the slots that must
 * start at zero itself).  Two calls that overlap on the GPU -- two host threads, or two streams -- race on it and both results
 * are garbage: serialise them. */
#define MPIC_KINETIC_STATIC_SCRATCH_BYTES 9728
#define MPIC_KINETIC_STATIC_MAX_PARTICLES 2147483584LL   /* 2^31 - 64 */
int mpic_kinetic_energy_static(const double* ux, const double* uy, const double* uz, const double* q, int64_t n, double qm,
                               double out[2], void* stream);

/* ---- tile ownership (owner[] map, PIC_2D.cpp:55,144-148; removeTileFromRank/addTileToRank, Auxiliar_functions.cpp:242-272).
 * On one GPU ownership is bookkeeping only (fields never change); it drives currentRank in the dumps. */
int mpic_get_owner(mpic_sim* s, int* owner /* global tiles */);
int mpic_set_owner(mpic_sim* s, const int* owner /* global tiles */);
int mpic_migrate_tile(mpic_sim* s, int gid, int dest_rank);

/* ---- run-time load balancing: cost-driven re-partition of the tile rows over the slabs, on the device.
 * Replaces the mechanism removeTileFromRank / addTileToRank (Auxiliar_functions.cpp:242-272: MPI_Send/Recv of one tile's grid, the
 * particle vector is dropped) and the hard-coded trigger of PIC_2D.cpp:399-412; the policy is the one the reference leaves as
 * future work (Report p.10).  A tile row is one contiguous slot range of every particle array (particles are tile-sorted) and
 * one contiguous row block of every field plane, so moving it is a handful of device-to-device copies over NVLink -- no host
 * staging, no new handle: the handle changes row0 / nrows IN PLACE.
 *
 * Policy (host, exact).  mpic_partition_rows: contiguous partition of n rows with the given costs into nranks non-empty slabs
 * minimising the most expensive slab (binary search on the bottleneck + greedy sweep); rows_out[2*k] = first row, [2*k+1] = rows.
 * mpic_rebalance_plan: cost of a tile row = its particles + w_cell * cells_per_tile_row (the per-cell work of the push and
 * the Yee sweeps does not vanish in empty tiles; w_cell ~ 8 on B200).  Returns 1 when new_rows differs from
 * old_rows AND lowers the bottleneck by more than min_gain (fraction), 0 when the partition should stay (new_rows = old_rows),
 * < 0 on error.  cost[0] / cost[1] (optional): bottleneck cost of the old / proposed partition. */
int mpic_partition_rows(int n, const double* row_cost, int nranks, int* rows_out);
int mpic_rebalance_plan(int global_tile_rows, int nranks, const int64_t* row_particles, int64_t cells_per_tile_row,
                        double w_cell, double min_gain, const int* old_rows, int* new_rows, double cost[2]);
/* Particles per LOCAL tile row (the occupancy the sort maintains), counts[nrows]. */
int mpic_get_tile_row_counts(mpic_sim* s, int64_t* counts);

/* Mechanism.  Collective over the ranks, between two steps (after mpic_sort_tiles / mpic_peer_step_end):
 *   1. every rank PACKS the global tile rows [row_lo, row_hi) of its slab that go to one new owner (itself included) into a
 *      parcel in device memory: the interior rows of the six field planes and the particles inside them (the current
 *      accumulators are zero between steps; guard rows are rebuilt).  The parcel is exported: mpic_parcel_info is an opaque,
 *      fixed-size, trivially copyable blob (CUDA IPC handle + sizes) that any transport can move (torch.distributed
 *      all_gather in slab.py; plain memory inside one process);
 *   2. every rank calls mpic_peer_disconnect; the blobs travel (this exchange is the barrier after which nobody maps anybody);
 *   3. every rank APPLIES: the handle is rebuilt in place on [new_row0, new_row0 + new_nrows) -- old device state freed, new
 *      state allocated, the parcels (which must tile the new rows exactly, any order) pulled over NVLink with device-to-device
 *      copies, x guards rebuilt, particles re-indexed (cell offsets; their order IS already the tile/cell order).
 *      new_max_particles == 0: keep the capacity if the new population fits, else grow to 1.25 x.  all_rows (optional,
 *      2*nranks ints: row0, nrows of every rank) updates the replicated owner[] map (PIC_2D.cpp:55) explicitly (no MPI_MAX
 *      trick);
 *   4. after a barrier of the caller's, mpic_reslab_finish frees the parcels this handle exported; the ring is re-connected
 *      (mpic_peer_export / _connect, or mpic_ring_enable_peer_access) and the y guards refreshed from the neighbours.
 * Results are partition-independent bit for bit (fixed-point current): a re-partitioned run equals the original. */
typedef struct mpic_parcel_info { unsigned char bytes[192]; } mpic_parcel_info;
int mpic_reslab_pack(mpic_sim* s, int row_lo, int row_hi, mpic_parcel_info* out);
int mpic_reslab_apply(mpic_sim* s, int new_row0, int new_nrows, int64_t new_max_particles, const mpic_parcel_info* parcels,
                      int nparcels, const int* all_rows);
int mpic_reslab_finish(mpic_sim* s);
/* rows of the handle as they are now */
int mpic_get_rows(mpic_sim* s, int* row0, int* nrows);

/* ---- single-process ring: one host thread drives N slab handles (sims[r] is rank r of N, one per GPU; handles may
 * share a device) and moves halo messages and migrating particles with CUDA peer copies over NVLink, ordered by events
 * -- no host synchronisation on the field path.  Replaces the MPI exchange phases (PIC_2D.cpp:200-248,276-319,348-396)
 * for the C++ driver; the one-process-per-GPU NCCL path (slab.py) uses the same kernels and gives the same bits. */
int mpic_ring_enable_peer_access(mpic_sim* const* sims, int n);
int mpic_ring_exchange(mpic_sim* const* sims, int n, int which);
int mpic_ring_refresh_guards(mpic_sim* const* sims, int n);          /* after uploads: y guards of E and B */
int mpic_ring_field_steps(mpic_sim* const* sims, int n, int nsteps);  /* phases A..F with ring exchanges */
int mpic_ring_migrate(mpic_sim* const* sims, int n);                  /* particle hand-off after a push */
int mpic_ring_pic_steps(mpic_sim* const* sims, int n, int nsteps);    /* full PIC steps */
/* The whole re-partition for a single-process ring: occupancy -> mpic_rebalance_plan -> pack / apply / finish on every handle
 * -> guards refreshed.  *changed = 1 when tile rows moved; rows_out (optional, 2*n ints) = the partition in force afterwards.
 * The trigger the reference hard-codes at step 5 (PIC_2D.cpp:399-412) becomes: call this every K steps (host/pic2d.cpp
 * --rebalance-every K). */
int mpic_ring_rebalance(mpic_sim* const* sims, int n, double w_cell, double min_gain, int* changed, int* rows_out);

/* ---- field dumps in the reference's HDF5 layout, written without libhdf5 (host code; works without a GPU).
 * mpic_save_rank_data replaces saveRankData (HDF5_output.h:11, HDF5_output.cpp:10-79): one file, for each listed tile a
 * group /Tile_<gid> with dataset `fields` (interior_ny+2g, interior_nx+2g) of the 48-byte compound {Ex,Ey,Ez,Bx,By,Bz}
 * (guards included) and int32 attributes tileRow, tileCol, currentRank.  grid6: ntiles patches, AoS, as Tile::grid. */
int mpic_save_rank_data(const char* filename, int ntiles, const int* gids, const int* tile_rows, const int* tile_cols,
                        const int* current_rank, int interior_nx, int interior_ny, int guard, const double* grid6);
/* Dump the tiles of this slab whose owner[] entry equals `rank` (all local tiles if rank < 0) to `filename`:
 * device -> tile patches with guards -> mpic_save_rank_data.  Phase H of the loop (PIC_2D.cpp:414-419). */
int mpic_dump_fields(mpic_sim* s, const char* filename, int rank);
/* The same dump, asynchronous: the tile patches are formed on the device and copied to pinned host memory on a side stream
 * behind the work queued so far; a writer thread then builds the HDF5 file while the caller keeps stepping (the step path
 * never waits for the disk).  At most one dump is in flight per handle: a second call first waits for the previous one.
 * mpic_dump_wait blocks until the file is complete and returns its status (0 when there is none). */
int mpic_dump_fields_async(mpic_sim* s, const char* filename, int rank);
int mpic_dump_wait(mpic_sim* s);
/* params.txt next to the dumps (PIC_2D.cpp:425-438), consumed by File_reader.py:read_params. */
int mpic_write_params(const char* path, double box_x, double box_y, int nx, int ny, int guard, int interior_nx,
                      int interior_ny, double sim_time, double dt, int total_steps);

/* ---- whole-step host-buffer entry point (the e2e path): upload fields+particles, nsteps PIC steps, download both. */
int mpic_pic_steps_host(mpic_sim* s, int nsteps, double* const planes[6], int64_t n, double* x, double* y, double* ux,
                        double* uy, double* uz, double* q, int* tag, int64_t* n_out);

/* ---- counters for the harness */
/* Device self-test: the push kernel's branch-free 1/sqrt(x) and a/b sequences against the compiler's IEEE sqrt/division
 * on n pseudo-random arguments of the range the push uses; *mismatches must come back 0.  (No reference counterpart.) */
int mpic_selftest_math(int64_t n, uint64_t seed, int64_t* mismatches);
int64_t mpic_kernel_launches(mpic_sim* s);   /* kernels launched by this handle so far */

#ifdef __cplusplus
}
#endif
#endif
