// ORACLE — test infrastructure only (see vo_math.h header).
#include "vo_math.h"
namespace vo {
float4x4 transpose(const float4x4 &a) {
    float4x4 r;
    for (int c = 0; c < 4; ++c) for (int k = 0; k < 4; ++k) r.m[c][k] = a.m[k][c];
    return r;
}
// Gauss-Jordan in double with partial pivoting; result rounded to fp32.
float4x4 inverse(const float4x4 &a) {
    double m[4][8];
    for (int r = 0; r < 4; ++r) {
        for (int c = 0; c < 4; ++c) { m[r][c] = a.m[c][r]; m[r][4 + c] = (r == c) ? 1.0 : 0.0; }
    }
    for (int i = 0; i < 4; ++i) {
        int p = i;
        for (int r = i + 1; r < 4; ++r) if (std::fabs(m[r][i]) > std::fabs(m[p][i])) p = r;
        if (p != i) for (int c = 0; c < 8; ++c) std::swap(m[i][c], m[p][c]);
        double d = m[i][i];
        if (d == 0.0) return float4x4::identity();
        for (int c = 0; c < 8; ++c) m[i][c] /= d;
        for (int r = 0; r < 4; ++r) if (r != i) {
            double f = m[r][i];
            if (f != 0.0) for (int c = 0; c < 8; ++c) m[r][c] -= f * m[i][c];
        }
    }
    float4x4 o;
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) o.m[c][r] = (float)m[r][4 + c];
    return o;
}
}  // namespace vo
