// compat.cpp -- the reference's C++ entry points (minipic_compat.h) implemented on the C ABI of libminipic_b200.so.
// Nothing here computes field values on the CPU: updateBhalf/updateEfull/packSendBuffer/updateGuardRegion each stage
// their host vectors through one CUDA launch (the L0 level of include/minipic_b200.h) and abort loudly if that fails.
#include "minipic_compat.h"
#include "../include/minipic_b200.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <filesystem>
#include <iostream>
#include <map>

namespace {
void must(int rc, const char* what) {
    if (rc < 0) {
        std::fprintf(stderr, "minipic host layer: %s failed (%d): %s\n", what, rc, mpic_last_error());
        std::abort();
    }
}
// default tile transport: in-process mailbox (single process hosting every rank, or a rank sending to itself)
std::map<int, std::deque<TileParcel>> g_mailbox;
void mailbox_send(int dest, const TileParcel& p) { g_mailbox[dest].push_back(p); }
bool mailbox_recv(int, int self, TileParcel& p) {
    auto& q = g_mailbox[self];
    if (q.empty()) return false;
    p = std::move(q.front());
    q.pop_front();
    return true;
}
tile_send_fn g_send = mailbox_send;
tile_recv_fn g_recv = mailbox_recv;
}  // namespace

void minipic_set_tile_transport(tile_send_fn send, tile_recv_fn recv) {
    g_send = send ? send : mailbox_send;
    g_recv = recv ? recv : mailbox_recv;
}

// ---- Field_update.h:36,54.  GridE/GridB vectors are AoS triplets == the ABI's E3/B3 arrays.
void updateBhalf(const vector<GridE>& Efull, const vector<GridB>& BhalfOld, vector<GridB>& BhalfNew, int interior_nx,
                 int interior_ny, int guard, double dt, double dx, double dy) {
    must(mpic_update_b_half(&Efull[0].Ex, &BhalfOld[0].Bx, &BhalfNew[0].Bx, interior_nx, interior_ny, guard, dt, dx, dy),
         "updateBhalf");
}
void updateEfull(vector<GridE>& Efull, const vector<GridB>& Bhalf, int interior_nx, int interior_ny, int guard, double dt,
                 double dx, double dy) {
    must(mpic_update_e_full(&Efull[0].Ex, &Bhalf[0].Bx, interior_nx, interior_ny, guard, dt, dx, dy), "updateEfull");
}

// ---- Auxiliar_functions.h:55-65 (pure integer; exact)
void findBestGrid(int size, int& R, int& C) { mpic_find_best_grid(size, &R, &C); }
void findBestTileGrid(int numTiles, int& tileRows, int& tileCols) { mpic_find_best_tile_grid(numTiles, &tileRows, &tileCols); }
int tileGlobalRow(int rankRow, int tileRows, int tRow) { return mpic_tile_global_row(rankRow, tileRows, tRow); }
int tileGlobalCol(int rankCol, int tileCols, int tCol) { return mpic_tile_global_col(rankCol, tileCols, tCol); }
int getGlobalID(int globalRow, int globalCol, int globalTileCols) { return mpic_get_global_id(globalRow, globalCol, globalTileCols); }
void getGlobalRowCol(int gID, int globalTileCols, int& row, int& col) { mpic_get_global_row_col(gID, globalTileCols, &row, &col); }
int getNeighborGID(int myGID, int d, int globalTileRows, int globalTileCols) {
    return mpic_get_neighbor_gid(myGID, d, globalTileRows, globalTileCols);
}
int computeTag(int tileGID, int direction) { return mpic_compute_tag(tileGID, direction); }

// ---- Auxiliar_functions.h:68-69
vector<Grid> packSendBuffer(const Tile& tile, int d, int guard) {
    vector<Grid> out((size_t)mpic_strip_cells(d, tile.info.nx, tile.info.ny, guard));
    must(mpic_pack_send_buffer(&tile.grid[0].Ex, tile.info.nx, tile.info.ny, guard, d, &out[0].Ex), "packSendBuffer");
    return out;
}
void updateGuardRegion(Tile& tile, int d, int guard, const vector<Grid>& rbuf) {
    must(mpic_update_guard_region(&tile.grid[0].Ex, tile.info.nx, tile.info.ny, guard, d, &rbuf[0].Ex), "updateGuardRegion");
}

// ---- Auxiliar_functions.h:72-73.  Same bookkeeping and console messages as the reference; the payload additionally
// carries the tile's particles (the reference drops them).
void removeTileFromRank(RankInfo& info, vector<int>& owner, int tileIndex, int dest) {
    TileParcel parcel;
    parcel.globalID = info.tiles[tileIndex].info.globalID;
    parcel.grid.swap(info.tiles[tileIndex].grid);
    parcel.particles.swap(info.tiles[tileIndex].particles);
    g_send(dest, parcel);
    owner[parcel.globalID] = dest;
    info.tiles.erase(info.tiles.begin() + tileIndex);
    std::cout << "Rank " << info.rank << " sent Tile GID=" << parcel.globalID << " to Rank " << dest << std::endl << std::flush;
}
void addTileToRank(RankInfo& info, vector<int>& owner, int source, int globalTileCols, int interior_nx, int interior_ny,
                   int guard) {
    TileParcel parcel;
    if (!g_recv(source, info.rank, parcel)) {
        std::fprintf(stderr, "minipic host layer: addTileToRank(rank %d <- %d): no tile in flight\n", info.rank, source);
        std::abort();
    }
    const size_t cells = (size_t)(interior_nx + 2 * guard) * (interior_ny + 2 * guard);
    if (parcel.grid.size() != cells) {
        std::fprintf(stderr, "minipic host layer: addTileToRank: tile %d has %zu cells, expected %zu\n", parcel.globalID,
                     parcel.grid.size(), cells);
        std::abort();
    }
    Tile t;
    t.info.globalID = parcel.globalID;
    getGlobalRowCol(parcel.globalID, globalTileCols, t.info.tileRow, t.info.tileCol);
    t.info.nx = interior_nx; t.info.ny = interior_ny; t.info.currentRank = info.rank;
    t.grid.swap(parcel.grid);
    t.particles.swap(parcel.particles);
    owner[parcel.globalID] = info.rank;
    info.tiles.push_back(std::move(t));
    std::cout << "Rank " << info.rank << " received Tile GID=" << parcel.globalID << " from Rank " << source << std::endl << std::flush;
}

// ---- Auxiliar_functions.h:75,78
void deleteFolder(const std::string& path) {
    std::error_code ec;
    std::filesystem::remove_all(path, ec);      // like the reference: silently nothing if the folder is absent
}
double heaviside(double x) { return x >= 0 ? 1.0 : 0.0; }

// ---- HDF5_output.h:11.  Like the reference (HDF5_output.cpp:70-78) a failed dump is reported and swallowed.
void saveRankData(const RankInfo& info, int guard, const std::string& filename) {
    const size_t n = info.tiles.size();
    vector<int> gid(n), row(n), col(n), cur(n);
    int tnx = n ? info.tiles[0].info.nx : 1, tny = n ? info.tiles[0].info.ny : 1;
    const size_t per = (size_t)(tnx + 2 * guard) * (tny + 2 * guard);
    vector<double> grid(per * 6 * n);
    for (size_t t = 0; t < n; t++) {
        const Tile& tile = info.tiles[t];
        gid[t] = tile.info.globalID; row[t] = tile.info.tileRow; col[t] = tile.info.tileCol; cur[t] = tile.info.currentRank;
        std::memcpy(&grid[t * per * 6], &tile.grid[0].Ex, per * 48);
    }
    if (mpic_save_rank_data(filename.c_str(), (int)n, gid.data(), row.data(), col.data(), cur.data(), tnx, tny, guard, grid.data()) < 0)
        std::fprintf(stderr, "saveRankData: %s\n", mpic_last_error());
}
