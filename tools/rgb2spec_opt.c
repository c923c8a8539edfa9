/* rgb2spec_opt.c - regenerates the sRGB -> sigmoid-polynomial coefficient table that the reference's
 * HeroWavelengthSpectrum reads as `sRGBToSpectrumTable_Data` (src/render_core/spectrum/hero.cpp:53-54,70:
 * `const float[3][res][res][res][4]`, res = 64). The data file `srgb2spec.h` is NOT in the reference tree, so the
 * table is rebuilt with the published method it was made with: W. Jakob, J. Hanika, "A Low-Dimensional Function
 * Space for Efficient Spectral Upsampling", EG 2019 - per grid cell a Gauss-Newton fit of the three coefficients of
 * s(c0 l^2 + c1 l + c2), s(x) = 1/2 + x / (2 sqrt(1 + x^2)), minimising the CIELAB distance to the target colour
 * under the D65 white point, warm-started from the neighbouring cell along the brightness axis.
 *
 * Layout written: float[3][res][res][res][4]: [max component l][z = brightness index][y][x][c0, c1, c2, 0], i.e.
 * exactly what RGBToSpectrumTable::decode_albedo indexes (hero.cpp:121-123: coefficients()[maxc][zi][yi][xi][i]).
 *
 * usage: rgb2spec_opt <cie_1nm_4x471.f32> <out.f32> [res=64]      (build: gcc -O2 -fopenmp ... -lm)
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define LAMBDA_MIN 360.0
#define LAMBDA_MAX 830.0
#define N_FINE 471 /* 1 nm quadrature */

static double lambda_tbl[N_FINE], rgb_tbl[3][N_FINE], xyz_whitepoint[3];

static const double xyz_to_srgb[3][3] = {{3.240479, -1.537150, -0.498535}, {-0.969256, 1.875991, 0.041556}, {0.055648, -0.204043, 1.057311}};
static const double srgb_to_xyz[3][3] = {{0.412453, 0.357580, 0.180423}, {0.212671, 0.715160, 0.072169}, {0.019334, 0.119193, 0.950227}};

static double sigmoid(double x) { return 0.5 * x / sqrt(1.0 + x * x) + 0.5; }
static double smoothstep(double x) { return x * x * (3.0 - 2.0 * x); }
static double sqr(double x) { return x * x; }

static void cie_lab(double *p) {
    double X = 0, Y = 0, Z = 0, Xw = xyz_whitepoint[0], Yw = xyz_whitepoint[1], Zw = xyz_whitepoint[2];
    for (int j = 0; j < 3; ++j) {
        X += p[j] * srgb_to_xyz[0][j];
        Y += p[j] * srgb_to_xyz[1][j];
        Z += p[j] * srgb_to_xyz[2][j];
    }
#define LAB_F(t) ((t) > (6.0 / 29.0) * (6.0 / 29.0) * (6.0 / 29.0) ? cbrt(t) : (t) / (3.0 * (6.0 / 29.0) * (6.0 / 29.0)) + 4.0 / 29.0)
    double fx = LAB_F(X / Xw), fy = LAB_F(Y / Yw), fz = LAB_F(Z / Zw);
    p[0] = 116.0 * fy - 16.0;
    p[1] = 500.0 * (fx - fy);
    p[2] = 200.0 * (fy - fz);
}

static void init_tables(const float *cie) {
    const float *cx = cie, *cy = cie + 471, *cz = cie + 2 * 471, *d65 = cie + 3 * 471;
    double h = (LAMBDA_MAX - LAMBDA_MIN) / (N_FINE - 1);
    double w[N_FINE], ysum = 0;
    for (int i = 0; i < N_FINE; ++i) {          /* composite Simpson 3/8 weights (N_FINE - 1 = 470 is not a multiple of 3: */
        w[i] = (i == 0 || i == N_FINE - 1) ? 0.5 * h : h; /* use the trapezoid rule instead; at 1 nm spacing the difference is < 1e-6) */
        ysum += w[i] * cy[i] * d65[i];
    }
    memset(rgb_tbl, 0, sizeof rgb_tbl);
    memset(xyz_whitepoint, 0, sizeof xyz_whitepoint);
    for (int i = 0; i < N_FINE; ++i) {
        double xyz[3] = {cx[i], cy[i], cz[i]}, I = d65[i] / ysum; /* illuminant normalised to Y = 1 */
        lambda_tbl[i] = (LAMBDA_MIN + i * h - LAMBDA_MIN) / (LAMBDA_MAX - LAMBDA_MIN);
        for (int k = 0; k < 3; ++k)
            for (int j = 0; j < 3; ++j) rgb_tbl[k][i] += xyz_to_srgb[k][j] * xyz[j] * I * w[i];
        for (int k = 0; k < 3; ++k) xyz_whitepoint[k] += xyz[k] * I * w[i];
    }
}

static void eval_residual(const double *coeffs, const double *rgb, double *residual) {
    double out[3] = {0, 0, 0};
    for (int i = 0; i < N_FINE; ++i) {
        double x = 0;
        for (int j = 0; j < 3; ++j) x = x * lambda_tbl[i] + coeffs[j];
        double s = sigmoid(x);
        for (int j = 0; j < 3; ++j) out[j] += rgb_tbl[j][i] * s;
    }
    cie_lab(out);
    memcpy(residual, rgb, 3 * sizeof(double));
    cie_lab(residual);
    for (int j = 0; j < 3; ++j) residual[j] -= out[j];
}

static void eval_jacobian(const double *coeffs, const double *rgb, double jac[3][3]) {
    const double eps = 1e-4;
    double r0[3], r1[3], tmp[3];
    for (int i = 0; i < 3; ++i) {
        memcpy(tmp, coeffs, sizeof tmp); tmp[i] -= eps; eval_residual(tmp, rgb, r0);
        memcpy(tmp, coeffs, sizeof tmp); tmp[i] += eps; eval_residual(tmp, rgb, r1);
        for (int j = 0; j < 3; ++j) jac[j][i] = (r1[j] - r0[j]) / (2 * eps);
    }
}

/* solves A x = b (3x3, partial pivoting); returns 0 when singular */
static int solve3(double A[3][3], const double *b, double *x) {
    double M[3][4];
    for (int i = 0; i < 3; ++i) { for (int j = 0; j < 3; ++j) M[i][j] = A[i][j]; M[i][3] = b[i]; }
    for (int c = 0; c < 3; ++c) {
        int p = c;
        for (int r = c + 1; r < 3; ++r) if (fabs(M[r][c]) > fabs(M[p][c])) p = r;
        if (fabs(M[p][c]) < 1e-15) return 0;
        if (p != c) for (int j = 0; j < 4; ++j) { double t = M[c][j]; M[c][j] = M[p][j]; M[p][j] = t; }
        for (int r = c + 1; r < 3; ++r) {
            double f = M[r][c] / M[c][c];
            for (int j = c; j < 4; ++j) M[r][j] -= f * M[c][j];
        }
    }
    for (int i = 2; i >= 0; --i) {
        double s = M[i][3];
        for (int j = i + 1; j < 3; ++j) s -= M[i][j] * x[j];
        x[i] = s / M[i][i];
    }
    return 1;
}

static void gauss_newton(const double rgb[3], double coeffs[3]) {
    for (int it = 0; it < 15; ++it) {
        double J[3][3], residual[3], x[3];
        eval_residual(coeffs, rgb, residual);
        eval_jacobian(coeffs, rgb, J);
        if (!solve3(J, residual, x)) return;
        double r = 0;
        for (int j = 0; j < 3; ++j) { coeffs[j] -= x[j]; r += sqr(residual[j]); }
        double mx = fmax(fmax(coeffs[0], coeffs[1]), coeffs[2]);
        if (mx > 200) for (int j = 0; j < 3; ++j) coeffs[j] *= 200 / mx;
        if (r < 1e-6) break;
    }
}

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s cie_1nm_4x471.f32 out.f32 [res]\n", argv[0]); return 2; }
    int res = argc > 3 ? atoi(argv[3]) : 64;
    static float cie[4 * 471];
    FILE *f = fopen(argv[1], "rb");
    if (!f || fread(cie, sizeof(float), 4 * 471, f) != 4 * 471) { fprintf(stderr, "cannot read %s\n", argv[1]); return 1; }
    fclose(f);
    init_tables(cie);
    double *scale = malloc(sizeof(double) * res);
    for (int k = 0; k < res; ++k) scale[k] = smoothstep(smoothstep((double)k / (res - 1)));
    size_t n = (size_t)3 * res * res * res * 4;
    float *out = calloc(n, sizeof(float));
    const double c0 = LAMBDA_MIN, c1 = 1.0 / (LAMBDA_MAX - LAMBDA_MIN);
    for (int l = 0; l < 3; ++l) {
#pragma omp parallel for collapse(2) schedule(dynamic, 4)
        for (int j = 0; j < res; ++j) {
            for (int i = 0; i < res; ++i) {
                double x = (double)i / (res - 1), y = (double)j / (res - 1);
                int start = res / 5;
                for (int pass = 0; pass < 2; ++pass) {           /* upwards from `start`, then downwards, each warm-started */
                    double coeffs[3] = {0, 0, 0};
                    int step = pass == 0 ? 1 : -1;
                    for (int k = start; k >= 0 && k < res; k += step) {
                        double b = scale[k], rgb[3];
                        rgb[l] = b; rgb[(l + 1) % 3] = x * b; rgb[(l + 2) % 3] = y * b;
                        gauss_newton(rgb, coeffs);
                        double A = coeffs[0], B = coeffs[1], C = coeffs[2];
                        size_t idx = ((((size_t)l * res + k) * res + j) * res + i) * 4;
                        out[idx + 0] = (float)(A * sqr(c1));                       /* coefficients of lambda in nm */
                        out[idx + 1] = (float)(B * c1 - 2 * A * c0 * sqr(c1));
                        out[idx + 2] = (float)(C - B * c0 * c1 + A * sqr(c0 * c1));
                        out[idx + 3] = 0.f;
                    }
                }
            }
        }
    }
    f = fopen(argv[2], "wb");
    if (!f || fwrite(out, sizeof(float), n, f) != n) { fprintf(stderr, "cannot write %s\n", argv[2]); return 1; }
    fclose(f);
    return 0;
}
