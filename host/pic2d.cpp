// pic2d.cpp -- the kept time-step driver (PIC_2D.cpp:22-463), parameterised and running on B200s through the C ABI.
//
// Same structure as the reference's main(): parameters -> decomposition -> initial condition -> owner map -> step-0 dump ->
// loop { phases A..F (half-B, E, half-B with their guard refreshes), G tile-migration demo, H dump } -> params.txt ->
// "Simulation complete.".  What changed: every literal of PIC_2D.cpp:36,58-74 is a command-line option whose default is
// the reference's value; the fields live on the GPU(s) for the whole run (L1 resident API) instead of being copied
// tile by tile around per-tile kernel calls; particles (which the reference declares but never implements) are optional.
// MPI ranks become (a) `--gpus N` slabs of whole tile rows, one handle per GPU, exchanged with peer copies (ring.cu), and
// (b) `--ranks R` logical owner ranks used only for the dump partition (one file per rank, currentRank attribute), so a
// run with --ranks 9 writes the same set of files and groups as the reference's 9-rank run.
#include "minipic_compat.h"
#include "../include/minipic_b200.h"
#include <sys/stat.h>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

namespace {

struct Options {
    int nx = 450, ny = 450;                 // PIC_2D.cpp:62
    double box_x = 10.0, box_y = 10.0;      // :58
    int guard = 2;                          // :65
    int ranks = 9;                          // the reference's documented run: mpirun -np 9
    int tiles_per_rank = 36;                // :36
    int tnx = 0, tny = 0;                   // override of interior_nx/ny (0: derive like PIC_2D.cpp:67-68)
    double sim_time = 500.0;                // :70
    double cfl = 0.5;                       // :72
    long steps = -1;                        // override of total_steps
    int save_frequency = 25;                // :74
    std::string ic = "test3";               // :104-141 (test1 | test2 | test3 | zero)
    std::string out = "Simulation/Fields";  // :151-163
    int gpus = 1;
    int scheme = 0, fma = 0;                // 1/1 replays the shipped dumps (old_files v1 loop, contracted build)
    int migrate_demo = 1;                   // phase G (:399-412) when ranks == 9
    int ppc = 0;                            // particles per cell (0: field solver only, like the reference)
    double u_th = 0.05, two_stream = 0.0, drift_x = 0.0;
    double beam_density = 0.0, beam_lo = 0.45, beam_hi = 0.55, beam_drift = 0.5, beam_uth = 0.01;
    int deposit = 0;                        // 0: charge-conserving zigzag; 1: direct CIC (SURVEY.md B.5 option a)
    unsigned long long seed = 12345;
    int diag_every = 0;
    int rebalance_every = 0;                // > 0: every K steps re-partition the tile rows over the GPUs when it pays (mpic_ring_rebalance)
    double rebalance_w_cell = 8.0, rebalance_min_gain = 0.05;
    bool async_dump = false;                // dumps leave the step path: D2H on a side stream + writer thread (mpic_dump_fields_async)
    bool quiet = false;
};

void usage() {
    std::puts("pic2d [--nx N --ny N --box-x L --box-y L --ranks R --tiles-per-rank T | --tile-nx n --tile-ny n]\n"
              "      [--sim-time t | --steps n] [--cfl f] [--save-frequency k] [--ic test1|test2|test3|zero] [--out dir]\n"
              "      [--gpus N] [--scheme 0|1 --fma 0|1] [--no-migrate-demo]\n"
              "      [--ppc n --uth u --two-stream u0 --drift-x u | --beam-density d --beam-lo f --beam-hi f --beam-drift u --beam-uth u]\n"
              "      [--deposit zigzag|cic] [--seed s] [--diag-every k] [--rebalance-every k --rebalance-w-cell w --rebalance-min-gain f] [--async-dump] [--quiet]\n"
              "defaults = the literals of the reference's PIC_2D.cpp (450x450, 9 ranks x 36 tiles of 25x25, Test-3 pulse, t=500)");
}

bool parse(int argc, char** argv, Options& o) {
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto val = [&]() -> const char* { if (i + 1 >= argc) { std::fprintf(stderr, "%s needs a value\n", a.c_str()); std::exit(2); } return argv[++i]; };
        if (a == "--nx") o.nx = atoi(val()); else if (a == "--ny") o.ny = atoi(val());
        else if (a == "--box-x") o.box_x = atof(val()); else if (a == "--box-y") o.box_y = atof(val());
        else if (a == "--ranks") o.ranks = atoi(val()); else if (a == "--tiles-per-rank") o.tiles_per_rank = atoi(val());
        else if (a == "--tile-nx") o.tnx = atoi(val()); else if (a == "--tile-ny") o.tny = atoi(val());
        else if (a == "--sim-time") o.sim_time = atof(val()); else if (a == "--steps") o.steps = atol(val());
        else if (a == "--cfl") o.cfl = atof(val()); else if (a == "--save-frequency") o.save_frequency = atoi(val());
        else if (a == "--ic") o.ic = val(); else if (a == "--out") o.out = val();
        else if (a == "--gpus") o.gpus = atoi(val()); else if (a == "--scheme") o.scheme = atoi(val()); else if (a == "--fma") o.fma = atoi(val());
        else if (a == "--no-migrate-demo") o.migrate_demo = 0;
        else if (a == "--ppc") o.ppc = atoi(val()); else if (a == "--uth") o.u_th = atof(val());
        else if (a == "--two-stream") o.two_stream = atof(val()); else if (a == "--drift-x") o.drift_x = atof(val());
        else if (a == "--beam-density") o.beam_density = atof(val()); else if (a == "--beam-lo") o.beam_lo = atof(val());
        else if (a == "--beam-hi") o.beam_hi = atof(val()); else if (a == "--beam-drift") o.beam_drift = atof(val());
        else if (a == "--beam-uth") o.beam_uth = atof(val());
        else if (a == "--deposit") { std::string v = val(); if (v != "zigzag" && v != "cic") { std::fprintf(stderr, "--deposit zigzag|cic\n"); return false; } o.deposit = v == "cic"; }
        else if (a == "--seed") o.seed = strtoull(val(), nullptr, 10); else if (a == "--diag-every") o.diag_every = atoi(val());
        else if (a == "--rebalance-every") o.rebalance_every = atoi(val()); else if (a == "--rebalance-w-cell") o.rebalance_w_cell = atof(val());
        else if (a == "--rebalance-min-gain") o.rebalance_min_gain = atof(val());
        else if (a == "--async-dump") o.async_dump = true;
        else if (a == "--quiet") o.quiet = true;
        else if (a == "--help" || a == "-h") { usage(); std::exit(0); }
        else { std::fprintf(stderr, "unknown option %s\n", a.c_str()); usage(); return false; }
    }
    return true;
}

void must(int rc, const char* what) {
    if (rc < 0) { std::fprintf(stderr, "pic2d: %s failed (%d): %s\n", what, rc, mpic_last_error()); std::exit(1); }
}

void mkdirs(const std::string& path) {
    std::string cur;
    for (size_t i = 0; i <= path.size(); i++) {
        if (i == path.size() || path[i] == '/') { if (!cur.empty()) mkdir(cur.c_str(), 0777); }
        if (i < path.size()) cur += path[i];
    }
}

// analytic initial conditions of the report's Tests 1-3 (PIC_2D.cpp:104-141), evaluated at the Yee stagger positions
// (Field_update.cpp:3-11).  Only interiors are filled: the guard ring is rebuilt on the device as the periodic image, which
// equals the reference's analytic guard fill for these box-periodic (or compactly supported) fields.
void initial_condition(const Options& o, double dx, double dy, std::vector<double> planes[6]) {
    const double A = 0.1, kx = 5 * 2 * M_PI / o.box_x, ky = 5 * 2 * M_PI / o.box_y, xc = 3.5, tau = 3.0;
    auto pulse = [&](double x) {
        const double c = std::cos((x - xc) / tau * M_PI * 0.5);
        return A * std::sin(kx * x) * c * c * heaviside(1 - std::fabs(x - xc) / tau);
    };
    for (int p = 0; p < 6; p++) planes[p].assign((size_t)o.nx * o.ny, 0.0);
    if (o.ic == "zero") return;
    for (int j = 0; j < o.ny; j++)
        for (int i = 0; i < o.nx; i++) {
            const size_t k = (size_t)j * o.nx + i;
            const double x_Ex = (i + 0.5) * dx, y_Ex = j * dy, x_Ey = i * dx, y_Ey = (j + 0.5) * dy, x_Bz = (i + 0.5) * dx, y_Bz = (j + 0.5) * dy;
            if (o.ic == "test1") { planes[0][k] = A * std::sin(ky * y_Ex); planes[5][k] = -A * std::sin(ky * y_Bz); }
            else if (o.ic == "test2") {
                planes[0][k] = (A / std::sqrt(2.0)) * std::sin(kx * x_Ex + ky * y_Ex);
                planes[1][k] = -(A / std::sqrt(2.0)) * std::sin(kx * x_Ey + ky * y_Ey);
                planes[5][k] = -A * std::sin(kx * x_Bz + ky * y_Bz);
            } else { planes[1][k] = pulse(x_Ey); planes[5][k] = pulse(x_Bz); }
        }
}

}  // namespace

int main(int argc, char** argv) {
    Options o;
    if (!parse(argc, argv, o)) return 2;

    // ---- decomposition, exactly as PIC_2D.cpp:29-54,67-68 (R x C ranks, tileRows x tileCols tiles per rank)
    int R, C, tileRows, tileCols;
    findBestGrid(o.ranks, R, C);
    findBestTileGrid(o.tiles_per_rank, tileRows, tileCols);
    int globalTileRows = R * tileRows, globalTileCols = C * tileCols;
    int interior_nx = o.nx / globalTileCols, interior_ny = o.ny / globalTileRows;
    if (o.tnx > 0 && o.tny > 0) {              // explicit tile shape: tile grid follows, ranks split it as evenly as findBestGrid allows
        interior_nx = o.tnx; interior_ny = o.tny;
        globalTileCols = o.nx / interior_nx; globalTileRows = o.ny / interior_ny;
        if (globalTileRows % R || globalTileCols % C) { R = o.ranks; C = 1; }
        if (globalTileRows % R) { std::fprintf(stderr, "pic2d: %d tile rows cannot be split over %d ranks\n", globalTileRows, o.ranks); return 2; }
        tileRows = globalTileRows / R; tileCols = globalTileCols / C;
    }
    if (interior_nx * globalTileCols != o.nx || interior_ny * globalTileRows != o.ny) {
        std::fprintf(stderr, "pic2d: nx, ny must be multiples of the tile grid (%d x %d tiles); the reference would silently drop cells (PIC_2D.cpp:67-68)\n",
                     globalTileCols, globalTileRows);
        return 2;
    }
    const int totalGlobalTiles = globalTileRows * globalTileCols;
    const double dx = o.box_x / o.nx, dy = o.box_y / o.ny;
    const double dt_courant = 1.0 / std::sqrt((1.0 / (dx * dx)) + (1.0 / (dy * dy)));
    const double dt = o.cfl * dt_courant;
    const long total_steps = o.steps >= 0 ? o.steps : static_cast<long>(o.sim_time / dt);
    if (globalTileRows < o.gpus) { std::fprintf(stderr, "pic2d: more GPUs than tile rows\n"); return 2; }

    // ---- owner[gid] = logical rank (PIC_2D.cpp:55,144-148)
    vector<int> owner(totalGlobalTiles, -1);
    for (int gr = 0; gr < globalTileRows; gr++)
        for (int gc = 0; gc < globalTileCols; gc++)
            owner[getGlobalID(gr, gc, globalTileCols)] = (gr / tileRows) * C + gc / tileCols;

    // ---- device handles: one slab of tile rows per GPU
    const int N = o.gpus;
    if (mpic_device_count() < 1) { std::fprintf(stderr, "pic2d: no CUDA device (there is no CPU fallback)\n"); return 1; }
    const int ndev = mpic_device_count();
    vector<mpic_sim*> sims(N, nullptr);
    long long max_particles = 0;
    vector<double> planes[6];
    initial_condition(o, dx, dy, planes);
    for (int r = 0; r < N; r++) {
        mpic_params p;
        std::memset(&p, 0, sizeof p);
        p.nx = o.nx; p.ny = o.ny; p.tnx = interior_nx; p.tny = interior_ny; p.guard = o.guard; p.scheme = o.scheme; p.fma_mode = o.fma;
        p.rank = r; p.nranks = N; p.device = r % ndev; p.dx = dx; p.dy = dy; p.dt = dt; p.qm = -1.0; p.deposit = o.deposit;
        mpic_slab_rows(globalTileRows, N, r, &p.row0, &p.nrows);
        if (o.ppc > 0) {
            const double cells = (double)o.nx * p.nrows * interior_ny;
            const double beam = o.beam_density > 0 ? o.ppc * o.beam_density * (double)o.nx * o.ny * (o.beam_hi - o.beam_lo) : 0.0;
            p.max_particles = (long long)(1.25 * (cells * o.ppc + beam)) + 65536;
            max_particles += p.max_particles;
        }
        must(mpic_create(&p, &sims[r]), "mpic_create");
        must(mpic_set_owner(sims[r], owner.data()), "mpic_set_owner");
        const double* ptr[6];
        for (int k = 0; k < 6; k++) ptr[k] = planes[k].data() + (size_t)p.row0 * interior_ny * o.nx;
        must(mpic_upload_planes(sims[r], ptr), "mpic_upload_planes");
        if (o.ppc > 0) {
            must(mpic_load_uniform(sims[r], o.ppc, o.u_th, o.drift_x, 0.0, 0.0, o.two_stream, 1.0, 0.0, 1.0, o.seed, 0), "mpic_load_uniform");
            if (o.beam_density > 0)      // dense drifting beam in a band of the domain (BASELINE.json configs[4])
                must(mpic_load_uniform(sims[r], (int)(o.ppc * o.beam_density), o.beam_uth, o.beam_drift, 0.0, 0.0, 0.0, o.beam_density,
                                       o.beam_lo, o.beam_hi, o.seed + 1, 1), "mpic_load_uniform(beam)");
        }
    }
    if (N > 1) {
        must(mpic_ring_enable_peer_access(sims.data(), N), "mpic_ring_enable_peer_access");
        must(mpic_ring_refresh_guards(sims.data(), N), "mpic_ring_refresh_guards");
    }

    // ---- output folder + step-0 dump (PIC_2D.cpp:151-168)
    deleteFolder(o.out);
    mkdirs(o.out);
    // which slabs hold tiles of which logical rank (a rank's tiles span several slabs when ranks do not nest in slabs)
    vector<vector<int>> tiles_on(o.ranks, vector<int>(N, 0));
    auto slab_of_row = [&](int gr) {                     // the rows a handle owns change when the load balancer runs
        for (int r = 0; r < N; r++) { int row0, nrows; mpic_get_rows(sims[r], &row0, &nrows); if (gr >= row0 && gr < row0 + nrows) return r; }
        return 0;
    };
    auto recount_tiles_on = [&]() {
        for (auto& v : tiles_on) std::fill(v.begin(), v.end(), 0);
        for (int gr = 0; gr < globalTileRows; gr++) {
            const int r = slab_of_row(gr);
            for (int gc = 0; gc < globalTileCols; gc++) {
                const int own = owner[getGlobalID(gr, gc, globalTileCols)];
                if (own >= 0 && own < o.ranks) tiles_on[own][r]++;
            }
        }
    };
    recount_tiles_on();
    auto dump = [&](long step) {
        for (int rank = 0; rank < o.ranks; rank++) {
            int holders = 0;
            for (int r = 0; r < N; r++) holders += tiles_on[rank][r] > 0;
            for (int r = 0; r < N; r++) {
                if (N > 1 && tiles_on[rank][r] == 0 && !(holders == 0 && r == 0)) continue;
                std::string fn = o.out + "/fields_rank_" + std::to_string(rank) + (holders > 1 ? "_gpu_" + std::to_string(r) : "") +
                                 "_step_" + std::to_string(step) + ".h5";
                if (o.async_dump && o.ranks == 1) must(mpic_dump_fields_async(sims[r], fn.c_str(), rank), "mpic_dump_fields_async");
                else must(mpic_dump_fields(sims[r], fn.c_str(), rank), "mpic_dump_fields");   // one dump in flight per handle: several
            }                                                                                 // logical ranks per slab stay synchronous
        }
    };
    auto diagnostics = [&](long step) {
        double tot[5] = { 0, 0, 0, 0, 0 };
        for (int r = 0; r < N; r++) {
            double d[5];
            must(mpic_diagnostics(sims[r], d), "mpic_diagnostics");
            tot[0] += d[0]; tot[1] += d[1]; tot[3] += d[3]; tot[4] += d[4];
            if (d[2] > tot[2]) tot[2] = d[2];
        }
        std::printf("step %ld t=%.6g field_energy=%.17g kinetic_energy=%.17g total=%.17g gauss_residual=%.3e charge=%.12g particles=%.0f\n",
                    step, step * dt, tot[0], tot[1], tot[0] + tot[1], N == 1 ? tot[2] : -1.0, tot[3], tot[4]);
    };
    if (o.save_frequency > 0) dump(0);
    if (o.diag_every > 0) diagnostics(0);

    // ---- time loop (PIC_2D.cpp:171-420).  Steps between two host-visible events are queued back to back on the device.
    auto t0 = std::chrono::steady_clock::now();
    long step = 0;
    while (step < total_steps) {
        long next = total_steps;
        if (o.save_frequency > 0) next = std::min(next, (step / o.save_frequency + 1) * (long)o.save_frequency);
        if (o.diag_every > 0) next = std::min(next, (step / o.diag_every + 1) * (long)o.diag_every);
        if (o.migrate_demo && o.ranks == 9 && step < 5) next = std::min(next, 5L);
        if (o.rebalance_every > 0 && N > 1) next = std::min(next, (step / o.rebalance_every + 1) * (long)o.rebalance_every);
        const int k = (int)(next - step);
        if (N == 1) must(o.ppc > 0 ? mpic_pic_steps(sims[0], k) : mpic_field_steps(sims[0], k), "time step");
        else must(o.ppc > 0 ? mpic_ring_pic_steps(sims.data(), N, k) : mpic_ring_field_steps(sims.data(), N, k), "time step (ring)");
        step = next;
        // --- Phase G: the reference's hard-coded migration demo (PIC_2D.cpp:399-412): at step 5 of a 9-rank run, ranks 0,1,2
        // each hand their local tile #3 to rank 3.  Field values never change; ownership and dump grouping do.
        if (o.migrate_demo && o.ranks == 9 && step == 5) {
            for (int src = 0; src < 3; src++) {
                const int gRow = tileGlobalRow(src / C, tileRows, 3 / tileCols), gCol = tileGlobalCol(src % C, tileCols, 3 % tileCols);
                const int gid = getGlobalID(gRow, gCol, globalTileCols);
                { const int r = slab_of_row(gRow); tiles_on[owner[gid]][r]--; tiles_on[3][r]++; }
                owner[gid] = 3;
                for (int r = 0; r < N; r++) must(mpic_migrate_tile(sims[r], gid, 3), "mpic_migrate_tile");
                if (!o.quiet) {
                    std::cout << "Rank " << src << " sent Tile GID=" << gid << " to Rank 3" << std::endl;
                    std::cout << "Rank 3 received Tile GID=" << gid << " from Rank " << src << std::endl;
                }
            }
        }
        // --- Phase G as a load balancer (the policy the report leaves as future work, p.10): every K steps the tile rows are
        // re-partitioned over the GPUs when the most loaded slab gets >= min_gain lighter; rows travel WITH their fields and
        // particles, device to device, and the handles change their rows in place
        if (o.rebalance_every > 0 && N > 1 && step % o.rebalance_every == 0 && step < total_steps) {
            int changed = 0;
            vector<int> rows(2 * (size_t)N);
            auto tb = std::chrono::steady_clock::now();
            must(mpic_ring_rebalance(sims.data(), N, o.rebalance_w_cell, o.rebalance_min_gain, &changed, rows.data()), "mpic_ring_rebalance");
            if (changed) {
                must(mpic_get_owner(sims[0], owner.data()), "mpic_get_owner");
                recount_tiles_on();
                if (!o.quiet) {
                    std::printf("step %ld: rebalanced in %.2f ms, tile rows per GPU:", step,
                                1e3 * std::chrono::duration<double>(std::chrono::steady_clock::now() - tb).count());
                    for (int r = 0; r < N; r++) std::printf(" %d", rows[2 * r + 1]);
                    std::printf("\n");
                }
            }
        }
        if (o.save_frequency > 0 && step % o.save_frequency == 0) dump(step);
        if (o.diag_every > 0 && step % o.diag_every == 0) diagnostics(step);
    }
    for (int r = 0; r < N; r++) { must(mpic_sync(sims[r]), "mpic_sync"); must(mpic_dump_wait(sims[r]), "mpic_dump_wait"); }
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    must(mpic_write_params((o.out + "/params.txt").c_str(), o.box_x, o.box_y, o.nx, o.ny, o.guard, interior_nx, interior_ny,
                           o.sim_time, dt, (int)total_steps), "mpic_write_params");
    if (!o.quiet) {
        long long launches = 0, np = 0;
        for (int r = 0; r < N; r++) { launches += mpic_kernel_launches(sims[r]); np += mpic_num_particles(sims[r]); }
        std::printf("%ld steps of %dx%d cells (%d x %d tiles of %dx%d) on %d GPU(s): %.3f s wall, %.3g cell-updates/s",
                    total_steps, o.nx, o.ny, globalTileCols, globalTileRows, interior_nx, interior_ny, N, secs,
                    secs > 0 ? (double)o.nx * o.ny * total_steps / secs : 0.0);
        if (np) std::printf(", %lld particles, %.3g particle-pushes/s", np, secs > 0 ? (double)np * total_steps / secs : 0.0);
        std::printf(" (includes dumps); %lld kernel launches\n", launches);
    }
    for (int r = 0; r < N; r++) mpic_destroy(sims[r]);
    std::cout << "\nSimulation complete.\n";
    return 0;
}
