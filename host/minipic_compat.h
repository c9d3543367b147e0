// minipic_compat.h -- the reference's C++ interface, re-exported on top of the C ABI (include/minipic_b200.h).
//
// A caller written against the reference's three headers (its only caller is main() in PIC_2D.cpp) compiles and links
// against host/ unchanged: same type names and layouts, same free-function names, argument order and meaning, same
// (absent) error behaviour -- the reference functions are void and never check; here a failing ABI call prints
// mpic_last_error() and aborts, because silently continuing with stale fields would be worse.
//   Field_update.h:9-17,36,54          GridE, GridB, updateBhalf, updateEfull        -> one CUDA launch per call (L0 compat)
//   Auxiliar_functions.h:11-49,55-78   tables, Particle, Grid, TileInfo, Tile, RankInfo, the 13 helpers
//   HDF5_output.h:11                   saveRankData                                   -> libhdf5-free writer
// The three forwarding headers next to this file carry the reference's header names.
#ifndef MINIPIC_COMPAT_H
#define MINIPIC_COMPAT_H
#include <cmath>
#include <string>
#include <vector>

using std::vector;

// ---- Field_update.h:9-17
struct GridE { double Ex, Ey, Ez; };
struct GridB { double Bx, By, Bz; };

// ---- Auxiliar_functions.h:11-13: neighbour directions 0 L, 1 R, 2 U(row-1), 3 D(row+1), 4 TL, 5 TR, 6 BL, 7 BR
static const int dRow[8] = { 0, 0, -1, 1, -1, -1, 1, 1 };
static const int dCol[8] = { -1, 1, 0, 0, -1, 1, -1, 1 };
static const int opposite[8] = { 1, 0, 3, 2, 7, 6, 5, 4 };

// ---- Auxiliar_functions.h:16-49 (layouts are part of the contract: Grid is the wire format and the HDF5 compound)
struct Particle { int tag; double q; double x, y; double px, py, pz; };
struct Grid { double Ex, Ey, Ez; double Bx, By, Bz; };
struct TileInfo { int globalID; int tileRow, tileCol; int nx, ny; int currentRank; };
struct Tile { TileInfo info; vector<Particle> particles; vector<Grid> grid; };
struct RankInfo { int rank, rankRow, rankCol; int tileRows, tileCols; vector<Tile> tiles; };
static_assert(sizeof(Grid) == 48 && sizeof(GridE) == 24 && sizeof(GridB) == 24 && sizeof(Particle) == 56, "reference POD layout");

// ---- Field_update.h:36,54
void updateBhalf(const vector<GridE>& Efull, const vector<GridB>& BhalfOld, vector<GridB>& BhalfNew, int interior_nx,
                 int interior_ny, int guard, double dt, double dx, double dy);
void updateEfull(vector<GridE>& Efull, const vector<GridB>& Bhalf, int interior_nx, int interior_ny, int guard, double dt,
                 double dx, double dy);

// ---- Auxiliar_functions.h:55-78
void findBestGrid(int size, int& R, int& C);
void findBestTileGrid(int numTiles, int& tileRows, int& tileCols);
int tileGlobalRow(int rankRow, int tileRows, int tRow);
int tileGlobalCol(int rankCol, int tileCols, int tCol);
int getGlobalID(int globalRow, int globalCol, int globalTileCols);
void getGlobalRowCol(int gID, int globalTileCols, int& row, int& col);
int getNeighborGID(int myGID, int d, int globalTileRows, int globalTileCols);
int computeTag(int tileGID, int direction);
vector<Grid> packSendBuffer(const Tile& tile, int d, int guard);
void updateGuardRegion(Tile& tile, int d, int guard, const vector<Grid>& rbuf);
void removeTileFromRank(RankInfo& info, vector<int>& owner, int tileIndex, int dest);
void addTileToRank(RankInfo& info, vector<int>& owner, int source, int globalTileCols, int interior_nx, int interior_ny,
                   int guard);
void deleteFolder(const std::string& path);
double heaviside(double x);

// ---- HDF5_output.h:11
void saveRankData(const RankInfo& info, int guard, const std::string& filename);

// ---- additions (not in the reference): the tile hand-off transport.  The reference moves a tile with MPI_Send/MPI_Recv
// (tags 999 and 1000, Auxiliar_functions.cpp:244-264).  Without MPI the two halves meet in an in-process mailbox keyed
// by destination rank; a multi-process host replaces it with minipic_set_tile_transport().  Unlike the reference
// (wart W5) the tile's particles travel with it.
struct TileParcel { int globalID; vector<Grid> grid; vector<Particle> particles; };
typedef void (*tile_send_fn)(int dest, const TileParcel& parcel);
typedef bool (*tile_recv_fn)(int source, int self, TileParcel& parcel);
void minipic_set_tile_transport(tile_send_fn send, tile_recv_fn recv);
#endif
