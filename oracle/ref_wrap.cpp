// oracle/ref_wrap.cpp -- TEST INFRASTRUCTURE ONLY.  Never linked into the product library.
//
// extern "C" wrappers that let Python (ctypes) call the REFERENCE's own, unmodified functions,
// compiled where they lie under /root/reference (see oracle/Makefile -> oracle/_ref/*.so):
//   updateBhalf / updateEfull                      Field_update.cpp:14-59 (or old_files/Field_update_v1.cpp)
//   packSendBuffer / updateGuardRegion             Auxiliar_functions.cpp:73-239
//   findBestGrid ... getNeighborGID / computeTag   Auxiliar_functions.cpp:16-70
//   removeTileFromRank / addTileToRank             Auxiliar_functions.cpp:242-272
// plus ref_run_loop(): a size-generic re-statement of the reference time loop's STRUCTURE
// (PIC_2D.cpp:171-420, phases A-G incl. its Grid<->GridE/GridB copies) that emulates all MPI ranks in
// one process.  The reference main() cannot be used directly because every size is a literal
// (PIC_2D.cpp:36,58-74).  No reference source text is copied here: only its functions are called.
#include <cstring>
#include <cstdio>
#include <deque>
#include <map>
#include <vector>
#include <chrono>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "Auxiliar_functions.h"   // reference header (needs shim mpi.h on the include path)
#include "Field_update.h"         // reference header

// ---------------------------------------------------------------- in-process MPI mailbox (shim/mpi.h)
namespace {
struct Msg { std::vector<unsigned char> bytes; };
std::map<int, std::deque<Msg>> g_mailbox;   // keyed by tag
bool g_quiet = true;
}
extern "C" int MPI_Send(const void* buf, int count, MPI_Datatype dt, int, int tag, MPI_Comm) {
    size_t n = (size_t)count * (dt == MPI_INT ? sizeof(int) : 1);
    Msg m; m.bytes.assign((const unsigned char*)buf, (const unsigned char*)buf + n);
    g_mailbox[tag].push_back(std::move(m));
    return 0;
}
extern "C" int MPI_Recv(void* buf, int count, MPI_Datatype dt, int, int tag, MPI_Comm, MPI_Status*) {
    size_t n = (size_t)count * (dt == MPI_INT ? sizeof(int) : 1);
    auto& q = g_mailbox[tag];
    if (q.empty() || q.front().bytes.size() != n) { std::fprintf(stderr, "ref_wrap: unmatched MPI_Recv tag %d\n", tag); return 1; }
    std::memcpy(buf, q.front().bytes.data(), n);
    q.pop_front();
    return 0;
}

static_assert(sizeof(Grid) == 48 && sizeof(GridE) == 24 && sizeof(GridB) == 24, "reference POD layout");

extern "C" {

int ref_sizeof_particle() { return (int)sizeof(Particle); }
#ifdef _OPENMP
int ref_num_threads() { return omp_get_max_threads(); }
void ref_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }   // a launcher (torchrun) may have exported OMP_NUM_THREADS=1
#else
int ref_num_threads() { return 1; }
void ref_set_threads(int) {}
#endif

// ---- Field_update.h:36,54 on raw pointers (AoS triplets, tile incl. guards, row-major j*TX+i)
void ref_update_b_half(const double* E3, const double* Bold3, double* Bnew3,
                       int tnx, int tny, int guard, double dt, double dx, double dy) {
    size_t n = (size_t)(tnx + 2 * guard) * (tny + 2 * guard);
    std::vector<GridE> E(n); std::vector<GridB> Bo(n), Bn(n);
    std::memcpy(E.data(), E3, n * 24); std::memcpy(Bo.data(), Bold3, n * 24); std::memcpy(Bn.data(), Bnew3, n * 24);
    updateBhalf(E, Bo, Bn, tnx, tny, guard, dt, dx, dy);
    std::memcpy(Bnew3, Bn.data(), n * 24);
}
void ref_update_e_full(double* E3, const double* B3, int tnx, int tny, int guard, double dt, double dx, double dy) {
    size_t n = (size_t)(tnx + 2 * guard) * (tny + 2 * guard);
    std::vector<GridE> E(n); std::vector<GridB> B(n);
    std::memcpy(E.data(), E3, n * 24); std::memcpy(B.data(), B3, n * 24);
    updateEfull(E, B, tnx, tny, guard, dt, dx, dy);
    std::memcpy(E3, E.data(), n * 24);
}

// ---- Auxiliar_functions.h:68-69
static void make_tile(Tile& t, const double* grid6, int tnx, int tny, int guard) {
    t.info.nx = tnx; t.info.ny = tny;
    size_t n = (size_t)(tnx + 2 * guard) * (tny + 2 * guard);
    t.grid.resize(n);
    if (grid6) std::memcpy(t.grid.data(), grid6, n * 48);
}
int ref_pack_send_buffer(const double* grid6, int tnx, int tny, int guard, int d, double* out6) {
    Tile t; make_tile(t, grid6, tnx, tny, guard);
    std::vector<Grid> b = packSendBuffer(t, d, guard);
    std::memcpy(out6, b.data(), b.size() * 48);
    return (int)b.size();
}
void ref_update_guard_region(double* grid6, int tnx, int tny, int guard, int d, const double* in6, int ncells) {
    Tile t; make_tile(t, grid6, tnx, tny, guard);
    std::vector<Grid> b(ncells); std::memcpy(b.data(), in6, (size_t)ncells * 48);
    updateGuardRegion(t, d, guard, b);
    std::memcpy(grid6, t.grid.data(), t.grid.size() * 48);
}

// ---- integer helpers, Auxiliar_functions.h:55-65
void ref_find_best_grid(int size, int* R, int* C) { findBestGrid(size, *R, *C); }
void ref_find_best_tile_grid(int n, int* r, int* c) { findBestTileGrid(n, *r, *c); }
int ref_tile_global_row(int rr, int tr, int t) { return tileGlobalRow(rr, tr, t); }
int ref_tile_global_col(int rc, int tc, int t) { return tileGlobalCol(rc, tc, t); }
int ref_get_global_id(int r, int c, int gtc) { return getGlobalID(r, c, gtc); }
void ref_get_global_row_col(int gid, int gtc, int* r, int* c) { getGlobalRowCol(gid, gtc, *r, *c); }
int ref_get_neighbor_gid(int gid, int d, int gtr, int gtc) { return getNeighborGID(gid, d, gtr, gtc); }
int ref_compute_tag(int gid, int d) { return computeTag(gid, d); }
double ref_heaviside(double x) { return heaviside(x); }
void ref_direction_tables(int* drow, int* dcol, int* opp) {
    for (int d = 0; d < 8; d++) { drow[d] = dRow[d]; dcol[d] = dCol[d]; opp[d] = opposite[d]; }
}

// ---------------------------------------------------------------------------------------------
// ref_run_loop: emulate `size = R*C` ranks x (tileRows x tileCols) tiles each, exactly the decomposition of
// PIC_2D.cpp:29-54,76-100, and run `steps` iterations of the loop structure PIC_2D.cpp:171-412.
//   global6   in/out: (ny, nx, 6) interior field, AoS {Ex,Ey,Ez,Bx,By,Bz}; guards are filled by periodic wrap
//             (equal to the analytic guard fill of PIC_2D.cpp:111-118 for any box-periodic IC).
//   phases    3 = current scheme (A half-B, C E, E half-B), 2 = old_files/PIC_2D_v1.cpp (A, C only).
//   mig       nmig x 4 ints {step, src_rank, local_tile_index, dest_rank}: generalises the literal demo at
//             PIC_2D.cpp:399-412 (step 5: ranks 0,1,2 each send local tile #3 to rank 3).
//   owner_out (globalTileRows*globalTileCols) final owner[] map; tiles_out (optional) per-gid tile patches
//             incl. guards (gid-major, TY*TX*6) as they would be dumped by saveRankData.
//   order_out (optional) for each rank, the gids in info.tiles order, rank-major, -1 padded to `totalTiles` each.
// Returns seconds spent in the step loop (wall clock) or <0 on error.
double ref_run_loop(double* global6, int nx, int ny, int R, int C, int tileRows, int tileCols, int guard,
                    double dt, double dx, double dy, int steps, int phases,
                    const int* mig, int nmig, int* owner_out, double* tiles_out, int* order_out) {
    const int size = R * C;
    const int gtr = R * tileRows, gtc = C * tileCols, total = gtr * gtc;
    if (nx % gtc || ny % gtr) return -1.0;                      // W7: the reference silently drops cells
    const int tnx = nx / gtc, tny = ny / gtr;
    const int TX = tnx + 2 * guard, TY = tny + 2 * guard;
    std::vector<int> owner(total, -1);
    std::vector<RankInfo> ranks(size);
    for (int r = 0; r < size; r++) {
        RankInfo& info = ranks[r];
        info.rank = r; info.rankRow = r / C; info.rankCol = r % C;
        info.tileRows = tileRows; info.tileCols = tileCols;
        info.tiles.resize(tileRows * tileCols);
        for (int t = 0; t < tileRows * tileCols; t++) {
            Tile& tile = info.tiles[t];
            int gRow = tileGlobalRow(info.rankRow, tileRows, t / tileCols);
            int gCol = tileGlobalCol(info.rankCol, tileCols, t % tileCols);
            tile.info.globalID = getGlobalID(gRow, gCol, gtc);
            tile.info.tileRow = gRow; tile.info.tileCol = gCol;
            tile.info.nx = tnx; tile.info.ny = tny; tile.info.currentRank = r;
            tile.grid.resize((size_t)TX * TY);
            for (int j = 0; j < TY; j++) for (int i = 0; i < TX; i++) {
                int gj = ((gRow * tny + j - guard) % ny + ny) % ny;
                int gi = ((gCol * tnx + i - guard) % nx + nx) % nx;
                std::memcpy(&tile.grid[(size_t)j * TX + i], global6 + ((size_t)gj * nx + gi) * 6, 48);
            }
            owner[tile.info.globalID] = r;
        }
    }
    // gid -> (rank, index) lookup rebuilt whenever ownership changes
    std::vector<std::pair<int,int>> where(total);
    auto rebuild = [&]() {
        for (int r = 0; r < size; r++) for (size_t i = 0; i < ranks[r].tiles.size(); i++)
            where[ranks[r].tiles[i].info.globalID] = { r, (int)i };
    };
    rebuild();

    // Flat list of every tile of every emulated rank.  The loops below are `omp parallel for` over it: ignored in libref.so
    // (built without -fopenmp: the reference's own single-threaded rank), active in libref_omp.so -- an all-cores variant of
    // the same loop, NOT the reference's parallelisation (that is MPI ranks), reported beside it as such.
    auto all_tiles = [&]() { std::vector<Tile*> v; for (auto& info : ranks) for (auto& t : info.tiles) v.push_back(&t); return v; };
    auto b_phase = [&]() {                                       // PIC_2D.cpp:174-196 / 322-344
        std::vector<Tile*> tl = all_tiles();
#pragma omp parallel for schedule(static)
        for (long ti = 0; ti < (long)tl.size(); ti++) {
            Tile& tile = *tl[ti];
            size_t n = tile.grid.size();
            std::vector<GridE> E(n); std::vector<GridB> Bo(n);
            for (size_t k = 0; k < n; k++) {
                E[k].Ex = tile.grid[k].Ex; E[k].Ey = tile.grid[k].Ey; E[k].Ez = tile.grid[k].Ez;
                Bo[k].Bx = tile.grid[k].Bx; Bo[k].By = tile.grid[k].By; Bo[k].Bz = tile.grid[k].Bz;
            }
            std::vector<GridB> Bn(n);
            updateBhalf(E, Bo, Bn, tnx, tny, guard, dt, dx, dy);
            for (size_t k = 0; k < n; k++) { tile.grid[k].Bx = Bn[k].Bx; tile.grid[k].By = Bn[k].By; tile.grid[k].Bz = Bn[k].Bz; }
        }
    };
    auto e_phase = [&]() {                                       // PIC_2D.cpp:251-272
        std::vector<Tile*> tl = all_tiles();
#pragma omp parallel for schedule(static)
        for (long ti = 0; ti < (long)tl.size(); ti++) {
            Tile& tile = *tl[ti];
            size_t n = tile.grid.size();
            std::vector<GridE> E(n); std::vector<GridB> B(n);
            for (size_t k = 0; k < n; k++) {
                E[k].Ex = tile.grid[k].Ex; E[k].Ey = tile.grid[k].Ey; E[k].Ez = tile.grid[k].Ez;
                B[k].Bx = tile.grid[k].Bx; B[k].By = tile.grid[k].By; B[k].Bz = tile.grid[k].Bz;
            }
            updateEfull(E, B, tnx, tny, guard, dt, dx, dy);
            for (size_t k = 0; k < n; k++) { tile.grid[k].Ex = E[k].Ex; tile.grid[k].Ey = E[k].Ey; tile.grid[k].Ez = E[k].Ez; }
        }
    };
    auto exchange = [&]() {                                      // PIC_2D.cpp:200-248: all sends packed first, then unpacked
        std::vector<std::vector<std::vector<Grid>>> rbuf(total, std::vector<std::vector<Grid>>(8));
        std::vector<Tile*> tl = all_tiles();
#pragma omp parallel for schedule(static)
        for (long ti = 0; ti < (long)tl.size(); ti++) {          // every (gid, d) writes its own mailbox slot
            Tile& tile = *tl[ti];
            int gid = tile.info.globalID;
            for (int d = 0; d < 8; d++) {
                int nbr = getNeighborGID(gid, d, gtr, gtc);
                // message tag computeTag(gid,d) is received by nbr as direction opposite[d]
                (void)owner[nbr]; (void)computeTag(gid, d);
                rbuf[nbr][opposite[d]] = packSendBuffer(tile, d, guard);
            }
        }
#pragma omp parallel for schedule(static)
        for (long ti = 0; ti < (long)tl.size(); ti++)
            for (int d = 0; d < 8; d++) updateGuardRegion(*tl[ti], d, guard, rbuf[tl[ti]->info.globalID][d]);
    };

    auto t0 = std::chrono::steady_clock::now();
    for (int step = 1; step <= steps; step++) {
        b_phase(); exchange();
        e_phase(); exchange();
        if (phases == 3) { b_phase(); exchange(); }
        bool moved = false;                                      // phase G, PIC_2D.cpp:399-412
        for (int m = 0; m < nmig; m++) if (mig[4 * m] == step) {
            int src = mig[4 * m + 1], idx = mig[4 * m + 2], dst = mig[4 * m + 3];
            if (src < 0 || src >= size || dst < 0 || dst >= size || idx >= (int)ranks[src].tiles.size()) return -2.0;
            removeTileFromRank(ranks[src], owner, idx, dst);
            addTileToRank(ranks[dst], owner, src, gtc, tnx, tny, guard);
            moved = true;
        }
        if (moved) rebuild();
    }
    double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    for (auto& info : ranks) for (auto& tile : info.tiles) {
        int gid = tile.info.globalID;
        for (int j = 0; j < tny; j++) for (int i = 0; i < tnx; i++)
            std::memcpy(global6 + ((size_t)(tile.info.tileRow * tny + j) * nx + tile.info.tileCol * tnx + i) * 6,
                        &tile.grid[(size_t)(j + guard) * TX + i + guard], 48);
        if (tiles_out) std::memcpy(tiles_out + (size_t)gid * TX * TY * 6, tile.grid.data(), (size_t)TX * TY * 48);
    }
    if (owner_out) std::memcpy(owner_out, owner.data(), total * sizeof(int));
    if (order_out) for (int r = 0; r < size; r++) for (int k = 0; k < total; k++)
        order_out[(size_t)r * total + k] = k < (int)ranks[r].tiles.size() ? ranks[r].tiles[k].info.globalID : -1;
    return secs;
}

} // extern "C"
