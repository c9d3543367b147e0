/* oracle/yee_oracle.c -- TEST INFRASTRUCTURE ONLY (checker; never shipped, never on the product path).
 *
 * Plain-C restatement of the reference's field-solver hot path:
 *   orc_update_b_half_tile / orc_update_e_full_tile   <- Field_update.cpp:14-36 / :40-59   (one tile incl. guards)
 *   orc_pack_send_buffer / orc_update_guard_region    <- Auxiliar_functions.cpp:73-169 / :172-239
 *   orc_find_best_grid ... orc_compute_tag            <- Auxiliar_functions.cpp:16-70
 *   orc_yee_steps_global                              <- the loop PIC_2D.cpp:171-396 collapsed onto ONE global periodic
 *                                                        array (legal because tiling+exchange is numerically a no-op,
 *                                                        SURVEY.md F8; pinned against oracle/_ref by tests/test_oracle_ref.py)
 * Pinned against: oracle/_ref (the reference's own sources compiled here) and the reference's shipped dumps
 * (tests/golden/pulse_rows.npz, made by tests/golden/make_golden.py from Simulation/Fields/*.h5).
 *
 * Arithmetic contract (shared, in words, with the CUDA kernels; see DESIGN.md "arithmetic contract"):
 *   mode 0 (STRICT): exactly the reference expression order, no contraction (build with -ffp-contract=off).
 *   mode 1 (FMA):    the contraction g++ -O2 -mfma -ffp-contract=fast applies to the reference expressions;
 *                    needed only to replay the shipped dumps bit-exactly (they were produced by an FMA build).
 *   scheme 0: current Field_update.cpp (B coefficient dt/(2*d)); scheme 1: old_files/Field_update_v1.cpp (dt/d).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ tile-level kernels (AoS triplets) */
void orc_update_b_half_tile(const double* E, const double* Bo, double* Bn, int tnx, int tny, int g,
                            double dt, double dx, double dy, int scheme, int mode) {
    int TX = tnx + 2 * g, TY = tny + 2 * g;
    double cx = scheme ? dt / dx : dt / (2 * dx), cy = scheme ? dt / dy : dt / (2 * dy);
    for (int j = g; j < TY - g; j++)
        for (int i = g; i < TX - g; i++) {
            const double* e = E + 3 * ((size_t)j * TX + i);
            const double* ex = e + 3;                 /* (i+1, j) */
            const double* ey = e + 3 * (size_t)TX;    /* (i, j+1) */
            const double* b = Bo + 3 * ((size_t)j * TX + i);
            double* o = Bn + 3 * ((size_t)j * TX + i);
            double dEz_y = ey[2] - e[2], dEz_x = ex[2] - e[2], dEy_x = ex[1] - e[1], dEx_y = ey[0] - e[0];
            if (mode == 0) {
                o[0] = b[0] - cy * dEz_y;
                o[1] = b[1] + cx * dEz_x;
                o[2] = b[2] - cx * dEy_x + cy * dEx_y;
            } else {
                o[0] = fma(-cy, dEz_y, b[0]);
                o[1] = fma(cx, dEz_x, b[1]);
                o[2] = fma(cy, dEx_y, fma(-cx, dEy_x, b[2]));
            }
        }
}

void orc_update_e_full_tile(double* E, const double* B, int tnx, int tny, int g,
                            double dt, double dx, double dy, int mode) {
    int TX = tnx + 2 * g, TY = tny + 2 * g;
    double cx = dt / dx, cy = dt / dy;
    for (int j = g; j < TY - g; j++)
        for (int i = g; i < TX - g; i++) {
            double* e = E + 3 * ((size_t)j * TX + i);
            const double* b = B + 3 * ((size_t)j * TX + i);
            const double* bx = b - 3;                 /* (i-1, j) */
            const double* by = b - 3 * (size_t)TX;    /* (i, j-1) */
            double dBz_y = b[2] - by[2], dBz_x = b[2] - bx[2], dBy_x = b[1] - bx[1], dBx_y = b[0] - by[0];
            if (mode == 0) {
                e[0] += cy * dBz_y;
                e[1] -= cx * dBz_x;
                e[2] += cx * dBy_x - cy * dBx_y;
            } else {
                e[0] = fma(cy, dBz_y, e[0]);
                e[1] = fma(-cx, dBz_x, e[1]);
                e[2] = e[2] + fma(cx, dBy_x, -(cy * dBx_y));
            }
        }
}

/* ------------------------------------------------------------------ guard strips (AoS Grid = 6 doubles) */
/* Strip geometry shared by pack (interior strip next to side d) and unpack (guard strip d).
 * d: 0 L, 1 R, 2 Up(-y), 3 Down(+y), 4 TL, 5 TR, 6 BL, 7 BR   (Auxiliar_functions.h:11-13) */
static void strip_geom(int d, int tnx, int tny, int g, int interior, int* r0, int* c0, int* nr, int* nc) {
    int colside = (d == 0 || d == 4 || d == 6) ? -1 : (d == 1 || d == 5 || d == 7) ? 1 : 0;
    int rowside = (d == 2 || d == 4 || d == 5) ? -1 : (d == 3 || d == 6 || d == 7) ? 1 : 0;
    if (colside == 0) { *c0 = g; *nc = tnx; }
    else { *nc = g; *c0 = interior ? (colside < 0 ? g : tnx) : (colside < 0 ? 0 : tnx + g); }
    if (rowside == 0) { *r0 = g; *nr = tny; }
    else { *nr = g; *r0 = interior ? (rowside < 0 ? g : tny) : (rowside < 0 ? 0 : tny + g); }
}
int orc_strip_cells(int d, int tnx, int tny, int g) { return d < 2 ? tny * g : d < 4 ? tnx * g : g * g; }

int orc_pack_send_buffer(const double* grid6, int tnx, int tny, int g, int d, double* out6) {
    int r0, c0, nr, nc, TX = tnx + 2 * g;
    strip_geom(d, tnx, tny, g, 1, &r0, &c0, &nr, &nc);
    for (int r = 0; r < nr; r++)
        memcpy(out6 + 6 * (size_t)r * nc, grid6 + 6 * ((size_t)(r0 + r) * TX + c0), (size_t)nc * 48);
    return nr * nc;
}
void orc_update_guard_region(double* grid6, int tnx, int tny, int g, int d, const double* in6) {
    int r0, c0, nr, nc, TX = tnx + 2 * g;
    strip_geom(d, tnx, tny, g, 0, &r0, &c0, &nr, &nc);
    for (int r = 0; r < nr; r++)
        memcpy(grid6 + 6 * ((size_t)(r0 + r) * TX + c0), in6 + 6 * (size_t)r * nc, (size_t)nc * 48);
}

/* ------------------------------------------------------------------ decomposition integers */
static const int k_drow[8] = { 0, 0, -1, 1, -1, -1, 1, 1 };
static const int k_dcol[8] = { -1, 1, 0, 0, -1, 1, -1, 1 };
static const int k_opp[8] = { 1, 0, 3, 2, 7, 6, 5, 4 };
void orc_direction_tables(int* drow, int* dcol, int* opp) {
    for (int d = 0; d < 8; d++) { drow[d] = k_drow[d]; dcol[d] = k_dcol[d]; opp[d] = k_opp[d]; }
}
void orc_find_best_grid(int size, int* R, int* C) {
    int r = (int)sqrt((double)size);
    while (r > 1 && size % r) r--;
    *R = r; *C = size / r;
}
int orc_get_global_id(int row, int col, int gtc) { return row * gtc + col; }
void orc_get_global_row_col(int gid, int gtc, int* row, int* col) { *row = gid / gtc; *col = gid % gtc; }
int orc_get_neighbor_gid(int gid, int d, int gtr, int gtc) {
    int row = gid / gtc + k_drow[d], col = gid % gtc + k_dcol[d];
    if (row < 0) row = gtr - 1; else if (row >= gtr) row = 0;
    if (col < 0) col = gtc - 1; else if (col >= gtc) col = 0;
    return row * gtc + col;
}
int orc_compute_tag(int gid, int d) { return gid * 8 + d; }

/* ------------------------------------------------------------------ global periodic planes (SoA, ny x nx each) */
#define WRAP(i, n) ((i) < 0 ? (i) + (n) : (i) >= (n) ? (i) - (n) : (i))

/* one B sweep: B <- B -/+ c*curl E with coefficient dt/(2d) (scheme 0) or dt/d (scheme 1) */
void orc_b_sweep_global(const double* ex, const double* ey, const double* ez, double* bx, double* by, double* bz,
                        int nx, int ny, double dt, double dx, double dy, int scheme, int mode) {
    double cx = scheme ? dt / dx : dt / (2 * dx), cy = scheme ? dt / dy : dt / (2 * dy);
    for (int j = 0; j < ny; j++) {
        int jp = WRAP(j + 1, ny);
        for (int i = 0; i < nx; i++) {
            int ip = WRAP(i + 1, nx);
            size_t k = (size_t)j * nx + i, kx = (size_t)j * nx + ip, ky = (size_t)jp * nx + i;
            double dEz_y = ez[ky] - ez[k], dEz_x = ez[kx] - ez[k], dEy_x = ey[kx] - ey[k], dEx_y = ex[ky] - ex[k];
            if (mode == 0) {
                bx[k] = bx[k] - cy * dEz_y;
                by[k] = by[k] + cx * dEz_x;
                bz[k] = bz[k] - cx * dEy_x + cy * dEx_y;
            } else {
                bx[k] = fma(-cy, dEz_y, bx[k]);
                by[k] = fma(cx, dEz_x, by[k]);
                bz[k] = fma(cy, dEx_y, fma(-cx, dEy_x, bz[k]));
            }
        }
    }
}
/* E sweep.  The in-place update is safe on a global array: E(k) depends on B only. */
void orc_e_sweep_global(double* ex, double* ey, double* ez, const double* bx, const double* by, const double* bz,
                        int nx, int ny, double dt, double dx, double dy, int mode) {
    double cx = dt / dx, cy = dt / dy;
    for (int j = 0; j < ny; j++) {
        int jm = WRAP(j - 1, ny);
        for (int i = 0; i < nx; i++) {
            int im = WRAP(i - 1, nx);
            size_t k = (size_t)j * nx + i, kx = (size_t)j * nx + im, ky = (size_t)jm * nx + i;
            double dBz_y = bz[k] - bz[ky], dBz_x = bz[k] - bz[kx], dBy_x = by[k] - by[kx], dBx_y = bx[k] - bx[ky];
            if (mode == 0) {
                ex[k] += cy * dBz_y;
                ey[k] -= cx * dBz_x;
                ez[k] += cx * dBy_x - cy * dBx_y;
            } else {
                ex[k] = fma(cy, dBz_y, ex[k]);
                ey[k] = fma(-cx, dBz_x, ey[k]);
                ez[k] = ez[k] + fma(cx, dBy_x, -(cy * dBx_y));
            }
        }
    }
}
/* The B sweep above updates B in place while reading E only, so it is also order-independent. */
void orc_yee_steps_global(double* ex, double* ey, double* ez, double* bx, double* by, double* bz,
                          int nx, int ny, double dt, double dx, double dy, int steps, int scheme, int mode) {
    for (int s = 0; s < steps; s++) {
        orc_b_sweep_global(ex, ey, ez, bx, by, bz, nx, ny, dt, dx, dy, scheme, mode);
        orc_e_sweep_global(ex, ey, ez, bx, by, bz, nx, ny, dt, dx, dy, mode);
        if (scheme == 0) orc_b_sweep_global(ex, ey, ez, bx, by, bz, nx, ny, dt, dx, dy, scheme, mode);
    }
}

/* discrete field energy 1/2 * sum dx dy (E^2 + B^2), SURVEY.md section 4 (v) */
double orc_field_energy(const double* ex, const double* ey, const double* ez, const double* bx, const double* by,
                        const double* bz, int nx, int ny, double dx, double dy) {
    double s = 0.0;
    for (size_t k = 0; k < (size_t)nx * ny; k++)
        s += ex[k] * ex[k] + ey[k] * ey[k] + ez[k] * ez[k] + bx[k] * bx[k] + by[k] * by[k] + bz[k] * bz[k];
    return 0.5 * s * dx * dy;
}
