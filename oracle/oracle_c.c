/* ORACLE / TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
 *
 * Plain-C restatement of the two floating-point-order-sensitive pieces of the
 * path, with the accumulation order spelled out in fma() calls:
 *   oracle_project      cvs_erosion.py:342-346 + kitti360scripts cam2image
 *   oracle_oriented_in  cvs_erosion.py:114-138
 * tests/test_oracle_golden.py checks that this scalar order reproduces numpy /
 * OpenBLAS bit for bit on the sample frames -- the claim the CUDA kernels'
 * __fma_rn chains rest on.  Build: make -C oracle  (gcc -O2 -ffp-contract=off).
 */
#include <math.h>
#include <stdint.h>

/* pts [n][4] float32, T 4x4 row-major, K 3x4 row-major -> u, v (int64), depth (f64) */
void oracle_project(const double* T, const double* K, const float* pts, int64_t n, int64_t* u, int64_t* v,
                    double* depth) {
    for (int64_t i = 0; i < n; ++i) {
        const double X = pts[4 * i], Y = pts[4 * i + 1], Z = pts[4 * i + 2];
        double p[3], q[3];
        for (int r = 0; r < 3; ++r) {
            double a = T[4 * r] * X;
            a = fma(T[4 * r + 1], Y, a);
            a = fma(T[4 * r + 2], Z, a);
            a = fma(T[4 * r + 3], 1.0, a);
            p[r] = a;
        }
        for (int r = 0; r < 3; ++r) {
            double a = K[4 * r] * p[0];
            a = fma(K[4 * r + 1], p[1], a);
            a = fma(K[4 * r + 2], p[2], a);
            q[r] = a;
        }
        double d = q[2];
        if (d == 0.0) d = -1e-6;
        const double ad = fabs(d);
        const double ru = rint(q[0] / ad), rv = rint(q[1] / ad);      /* np.round: half to even */
        /* numpy float64 -> int64 (cvttsd2si): NaN / out of range -> INT64_MIN */
        u[i] = (fabs(ru) < 9.2233720368547758e18) ? (int64_t)ru : INT64_MIN;
        v[i] = (fabs(rv) < 9.2233720368547758e18) ? (int64_t)rv : INT64_MIN;
        depth[i] = d;
    }
}

/* points [n][3] float32, corners [8][3] f64 -> inside[n] (0/1), closed slabs along c1-c0, c3-c0, c4-c0 */
void oracle_oriented_in(const float* pts, int64_t n, const double* c, uint8_t* inside) {
    const int idx[3] = {1, 3, 4};
    double e[3][3], ee[3];
    for (int a = 0; a < 3; ++a) {
        for (int k = 0; k < 3; ++k) e[a][k] = c[3 * idx[a] + k] - c[k];
        ee[a] = fma(e[a][2], e[a][2], fma(e[a][1], e[a][1], e[a][0] * e[a][0]));
    }
    for (int64_t i = 0; i < n; ++i) {
        const double rx = (double)pts[3 * i] - c[0], ry = (double)pts[3 * i + 1] - c[1],
                     rz = (double)pts[3 * i + 2] - c[2];
        int in = 1;
        for (int a = 0; a < 3; ++a) {
            const double d = fma(rz, e[a][2], fma(ry, e[a][1], rx * e[a][0]));
            const double t = d / ee[a];
            if (!(t >= 0.0 && t <= 1.0)) in = 0;
        }
        inside[i] = (uint8_t)in;
    }
}
