// TEST INFRASTRUCTURE — CPU restatement of the reference rasterizer (never linked into the product).
//
// The reference rasterizer (submodules/depth-diff-gaussian-rasterization, "RAST/") has no CPU
// implementation; this file restates its algorithm so that GPU-less CI can check the host logic and
// so that the CUDA kernels have a second, independent checker.  Every function cites the reference
// lines it follows.  Two scalar types are instantiated:
//   Ff     — FP32 with EMULATED nvcc FMA contraction (see below): meant to reproduce the reference's
//            rounding sequence for radii / depth bits / tile rects;
//   double — FP64 arbiter for tolerance-based comparisons (images, gradients).
//
// FMA emulation: nvcc (-fmad=true, the reference's default build, RAST/setup.py:29) contracts
// fadd(fmul(a,b), c) -> fma(a,b,c), preferring the FIRST operand when both are products (observed
// in the PTX of the reference's preprocessCUDA: `mul t, y, m4; fma t, x, m0, t; fma t, z, m8, t;
// add t, m12, t` for m0*x + m4*y + m8*z + m12).  Ff carries "I am an unrounded-able product"
// operands so that operator+/- can apply the same rule.  Compile with -ffp-contract=off.
//
// Pinning: tests/golden/raster_*.npz hold outputs of the REAL reference CUDA build (oracle/_ref,
// run on the GPU box by tests/golden/make_raster_golden.py); tests check this file against them.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace {

struct Ff {
    float v; bool m; float a, b;
    Ff() : v(0.f), m(false), a(0.f), b(0.f) {}
    Ff(float x) : v(x), m(false), a(0.f), b(0.f) {}
    Ff(double x) : v((float)x), m(false), a(0.f), b(0.f) {}
    Ff(int x) : v((float)x), m(false), a(0.f), b(0.f) {}
};
inline Ff operator*(Ff x, Ff y) { Ff r; r.v = x.v * y.v; r.m = true; r.a = x.v; r.b = y.v; return r; }
inline Ff operator+(Ff x, Ff y) {
    if (x.m) return Ff(std::fmaf(x.a, x.b, y.v));
    if (y.m) return Ff(std::fmaf(y.a, y.b, x.v));
    return Ff(x.v + y.v);
}
inline Ff operator-(Ff x, Ff y) {
    if (x.m) return Ff(std::fmaf(x.a, x.b, -y.v));
    if (y.m) return Ff(std::fmaf(-y.a, y.b, x.v));
    return Ff(x.v - y.v);
}
inline Ff operator-(Ff x) { Ff r = x; r.v = -x.v; if (x.m) r.a = -x.a; return r; }
inline Ff operator/(Ff x, Ff y) { return Ff(x.v / y.v); }
inline Ff& operator+=(Ff& x, Ff y) { x = x + y; return x; }
inline Ff& operator*=(Ff& x, Ff y) { x = x * y; return x; }
inline bool operator<(Ff x, Ff y) { return x.v < y.v; }
inline bool operator>(Ff x, Ff y) { return x.v > y.v; }
inline bool operator<=(Ff x, Ff y) { return x.v <= y.v; }
inline bool operator>=(Ff x, Ff y) { return x.v >= y.v; }
inline bool operator==(Ff x, Ff y) { return x.v == y.v; }
inline bool operator!=(Ff x, Ff y) { return x.v != y.v; }

template <class S> struct Ops;
template <> struct Ops<Ff> {
    static Ff sqrt(Ff x) { return Ff(std::sqrt(x.v)); }
    static Ff exp(Ff x) { return Ff(std::exp(x.v)); }
    static Ff ceil(Ff x) { return Ff(std::ceil(x.v)); }
    static Ff min(Ff x, Ff y) { return Ff(std::fmin(x.v, y.v)); }
    static Ff max(Ff x, Ff y) { return Ff(std::fmax(x.v, y.v)); }
    static float f(Ff x) { return x.v; }
    static double d(Ff x) { return (double)x.v; }
    static Ff lit(double x) { return Ff((float)x); }
    // ndc2Pix is evaluated in double in the reference (auxiliary.h:41-44: 1.0 / 0.5 literals)
    static Ff ndc2pix(Ff v, int S_) { return Ff((float)((((double)v.v + 1.0) * (double)S_ - 1.0) * 0.5)); }
};
template <> struct Ops<double> {
    static double sqrt(double x) { return std::sqrt(x); }
    static double exp(double x) { return std::exp(x); }
    static double ceil(double x) { return std::ceil(x); }
    static double min(double x, double y) { return std::fmin(x, y); }
    static double max(double x, double y) { return std::fmax(x, y); }
    static float f(double x) { return (float)x; }
    static double d(double x) { return x; }
    static double lit(double x) { return x; }
    static double ndc2pix(double v, int S_) { return ((v + 1.0) * S_ - 1.0) * 0.5; }
};

// glm-style column-major 3x3: m[c][r]; product in glm's summation order
// (third_party/glm/glm/detail/type_mat3x3.inl operator*: SrcA0r*SrcBc0 + SrcA1r*SrcBc1 + SrcA2r*SrcBc2).
template <class S> struct Mat3 { S m[3][3]; };
template <class S> Mat3<S> mul(const Mat3<S>& A, const Mat3<S>& B) {
    Mat3<S> R;
    for (int c = 0; c < 3; ++c)
        for (int r = 0; r < 3; ++r) R.m[c][r] = A.m[0][r] * B.m[c][0] + A.m[1][r] * B.m[c][1] + A.m[2][r] * B.m[c][2];
    return R;
}
template <class S> Mat3<S> transpose(const Mat3<S>& A) {
    Mat3<S> R;
    for (int c = 0; c < 3; ++c)
        for (int r = 0; r < 3; ++r) R.m[c][r] = A.m[r][c];
    return R;
}

const double SH_C0 = 0.28209479177387814, SH_C1 = 0.4886025119029199;
const double SH_C2[5] = {1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396};
const double SH_C3[7] = {-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
                         -0.4570457994644658, 1.445305721320277, -0.5900435899266435};

struct Scene {
    int P, D, M, W, H;
    float tanfovx, tanfovy, scale_modifier;
    const float *bg, *means3D, *shs, *colors_precomp, *opacities, *scales, *rotations, *cov3D_precomp, *view, *proj, *campos;
};

template <class S> struct Geom {
    std::vector<S> depth, mx, my, cov3D, ca, cb, cc, op, rgb;
    std::vector<int> radii;
    std::vector<uint8_t> clamped;     // 3 per Gaussian
    std::vector<uint32_t> tiles;
};

// transformPoint4x3 / 4x4 (auxiliary.h:58-77)
template <class S> void xf43(const S p[3], const float* m, S out[3]) {
    for (int i = 0; i < 3; ++i) out[i] = S(m[i]) * p[0] + S(m[4 + i]) * p[1] + S(m[8 + i]) * p[2] + S(m[12 + i]);
}
template <class S> void xf44(const S p[3], const float* m, S out[4]) {
    for (int i = 0; i < 4; ++i) out[i] = S(m[i]) * p[0] + S(m[4 + i]) * p[1] + S(m[8 + i]) * p[2] + S(m[12 + i]);
}

// getRect (auxiliary.h:46-56); p and radius -> tile rect
template <class S> void get_rect(S px, S py, int radius, int gx, int gy, int& x0, int& y0, int& x1, int& y1) {
    typedef Ops<S> O;
    const S r = S(radius), B = S(16);
    auto toi = [](S v) { double d = O::d(v); if (d > 2147483647.0) return 2147483647; if (d < -2147483648.0) return (int)-2147483647 - 1; return (int)d; };
    x0 = std::min(gx, std::max(0, toi((px - r) / B)));
    y0 = std::min(gy, std::max(0, toi((py - r) / B)));
    x1 = std::min(gx, std::max(0, toi((((px + r) + B) - S(1)) / B)));
    y1 = std::min(gy, std::max(0, toi((((py + r) + B) - S(1)) / B)));
}

// computeCov3D (forward.cu:118-152)
template <class S> void cov3d(const S s[3], S mod, const S q[4], S out[6], Mat3<S>* Rout = nullptr, Mat3<S>* Mout = nullptr) {
    Mat3<S> Sm;
    for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) Sm.m[c][r] = S(c == r ? 1 : 0);
    Sm.m[0][0] = mod * s[0]; Sm.m[1][1] = mod * s[1]; Sm.m[2][2] = mod * s[2];
    const S r = q[0], x = q[1], y = q[2], z = q[3];
    const S one = S(1), two = S(2);
    Mat3<S> R;
    R.m[0][0] = one - two * (y * y + z * z); R.m[0][1] = two * (x * y - r * z);       R.m[0][2] = two * (x * z + r * y);
    R.m[1][0] = two * (x * y + r * z);       R.m[1][1] = one - two * (x * x + z * z); R.m[1][2] = two * (y * z - r * x);
    R.m[2][0] = two * (x * z - r * y);       R.m[2][1] = two * (y * z + r * x);       R.m[2][2] = one - two * (x * x + y * y);
    Mat3<S> M = mul(Sm, R);
    Mat3<S> Sg = mul(transpose(M), M);
    out[0] = Sg.m[0][0]; out[1] = Sg.m[0][1]; out[2] = Sg.m[0][2]; out[3] = Sg.m[1][1]; out[4] = Sg.m[1][2]; out[5] = Sg.m[2][2];
    if (Rout) *Rout = R;
    if (Mout) *Mout = M;
}

// computeCov2D (forward.cu:74-113); also returns T, clamped t and the clamp flags for the backward
template <class S> void cov2d(const S mean[3], S fx, S fy, S tanx, S tany, const S c6[6], const float* V, S cov[3],
                              Mat3<S>* Tout = nullptr, S* tout = nullptr, S* xmul = nullptr, S* ymul = nullptr) {
    typedef Ops<S> O;
    S t[3];
    xf43(mean, V, t);
    const S limx = O::lit(1.3f) * tanx, limy = O::lit(1.3f) * tany;
    const S txtz = t[0] / t[2], tytz = t[1] / t[2];
    t[0] = O::min(limx, O::max(-limx, txtz)) * t[2];
    t[1] = O::min(limy, O::max(-limy, tytz)) * t[2];
    if (xmul) *xmul = (txtz < -limx || txtz > limx) ? S(0) : S(1);
    if (ymul) *ymul = (tytz < -limy || tytz > limy) ? S(0) : S(1);
    Mat3<S> J;
    J.m[0][0] = fx / t[2]; J.m[0][1] = S(0);      J.m[0][2] = -(fx * t[0]) / (t[2] * t[2]);
    J.m[1][0] = S(0);      J.m[1][1] = fy / t[2]; J.m[1][2] = -(fy * t[1]) / (t[2] * t[2]);
    J.m[2][0] = S(0);      J.m[2][1] = S(0);      J.m[2][2] = S(0);
    Mat3<S> Wm;
    Wm.m[0][0] = S(V[0]); Wm.m[0][1] = S(V[4]); Wm.m[0][2] = S(V[8]);
    Wm.m[1][0] = S(V[1]); Wm.m[1][1] = S(V[5]); Wm.m[1][2] = S(V[9]);
    Wm.m[2][0] = S(V[2]); Wm.m[2][1] = S(V[6]); Wm.m[2][2] = S(V[10]);
    Mat3<S> T = mul(Wm, J);
    Mat3<S> Vrk;
    Vrk.m[0][0] = c6[0]; Vrk.m[0][1] = c6[1]; Vrk.m[0][2] = c6[2];
    Vrk.m[1][0] = c6[1]; Vrk.m[1][1] = c6[3]; Vrk.m[1][2] = c6[4];
    Vrk.m[2][0] = c6[2]; Vrk.m[2][1] = c6[4]; Vrk.m[2][2] = c6[5];
    Mat3<S> cv = mul(mul(transpose(T), transpose(Vrk)), T);
    cov[0] = cv.m[0][0] + O::lit(0.3f);
    cov[1] = cv.m[0][1];
    cov[2] = cv.m[1][1] + O::lit(0.3f);
    if (Tout) *Tout = T;
    if (tout) { tout[0] = t[0]; tout[1] = t[1]; tout[2] = t[2]; }
}

// computeColorFromSH forward (forward.cu:20-71)
template <class S> void sh_color(int deg, int M, const S pos[3], const float* campos, const float* sh /*[M][3]*/, S out[3], uint8_t cl[3]) {
    typedef Ops<S> O;
    S dir[3] = {pos[0] - S(campos[0]), pos[1] - S(campos[1]), pos[2] - S(campos[2])};
    const S len = O::sqrt(dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2]);
    const S x = dir[0] / len, y = dir[1] / len, z = dir[2] / len;
    for (int c = 0; c < 3; ++c) {
        auto s = [&](int k) { return S(sh[3 * k + c]); };
        S v = O::lit((float)SH_C0) * s(0);
        if (deg > 0) {
            const S c1 = O::lit((float)SH_C1);
            v = v - c1 * y * s(1) + c1 * z * s(2) - c1 * x * s(3);
            if (deg > 1) {
                const S xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                v = v + O::lit((float)SH_C2[0]) * xy * s(4) + O::lit((float)SH_C2[1]) * yz * s(5) +
                    O::lit((float)SH_C2[2]) * (S(2) * zz - xx - yy) * s(6) + O::lit((float)SH_C2[3]) * xz * s(7) +
                    O::lit((float)SH_C2[4]) * (xx - yy) * s(8);
                if (deg > 2) {
                    v = v + O::lit((float)SH_C3[0]) * y * (S(3) * xx - yy) * s(9) + O::lit((float)SH_C3[1]) * xy * z * s(10) +
                        O::lit((float)SH_C3[2]) * y * (S(4) * zz - xx - yy) * s(11) +
                        O::lit((float)SH_C3[3]) * z * (S(2) * zz - S(3) * xx - S(3) * yy) * s(12) +
                        O::lit((float)SH_C3[4]) * x * (S(4) * zz - xx - yy) * s(13) +
                        O::lit((float)SH_C3[5]) * z * (xx - yy) * s(14) + O::lit((float)SH_C3[6]) * x * (xx - S(3) * yy) * s(15);
                }
            }
        }
        v = v + O::lit(0.5f);
        cl[c] = (v < S(0)) ? 1 : 0;
        out[c] = O::max(v, S(0));
    }
    (void)M;
}

// preprocessCUDA forward (forward.cu:155-256)
template <class S> void preprocess(const Scene& sc, Geom<S>& g) {
    typedef Ops<S> O;
    const int P = sc.P;
    const int gx = (sc.W + 15) / 16, gy = (sc.H + 15) / 16;
    const S focal_y = S(sc.H / (2.0f * sc.tanfovy)), focal_x = S(sc.W / (2.0f * sc.tanfovx));   // rasterizer_impl.cu:223-224 (host float)
    g.depth.assign(P, S(0)); g.mx.assign(P, S(0)); g.my.assign(P, S(0)); g.cov3D.assign(6 * (size_t)P, S(0));
    g.ca.assign(P, S(0)); g.cb.assign(P, S(0)); g.cc.assign(P, S(0)); g.op.assign(P, S(0)); g.rgb.assign(3 * (size_t)P, S(0));
    g.radii.assign(P, 0); g.clamped.assign(3 * (size_t)P, 0); g.tiles.assign(P, 0);
    for (int i = 0; i < P; ++i) {
        const S p[3] = {S(sc.means3D[3 * i]), S(sc.means3D[3 * i + 1]), S(sc.means3D[3 * i + 2])};
        S pv[3];
        xf43(p, sc.view, pv);
        if (pv[2] <= O::lit(0.2f)) continue;                      // in_frustum (auxiliary.h:139-164)
        S ph[4];
        xf44(p, sc.proj, ph);
        const S pw = S(1) / (ph[3] + O::lit(0.0000001f));
        const S ppx = ph[0] * pw, ppy = ph[1] * pw;
        S c6[6];
        if (sc.cov3D_precomp) {
            for (int k = 0; k < 6; ++k) c6[k] = S(sc.cov3D_precomp[6 * (size_t)i + k]);
        } else {
            const S s3[3] = {S(sc.scales[3 * i]), S(sc.scales[3 * i + 1]), S(sc.scales[3 * i + 2])};
            const S q4[4] = {S(sc.rotations[4 * i]), S(sc.rotations[4 * i + 1]), S(sc.rotations[4 * i + 2]), S(sc.rotations[4 * i + 3])};
            cov3d(s3, S(sc.scale_modifier), q4, c6);
            for (int k = 0; k < 6; ++k) { c6[k] = S(O::d(c6[k])); g.cov3D[6 * (size_t)i + k] = c6[k]; }
        }
        S cov[3];
        cov2d(p, focal_x, focal_y, S(sc.tanfovx), S(sc.tanfovy), c6, sc.view, cov);
        for (int k = 0; k < 3; ++k) cov[k] = S(O::d(cov[k]));    // returned as a float3: plain values from here on
        const S det = cov[0] * cov[2] - cov[1] * cov[1];
        if (det == S(0)) continue;
        const S det_inv = S(1) / det;
        const S conic[3] = {cov[2] * det_inv, -cov[1] * det_inv, cov[0] * det_inv};
        const S mid = O::lit(0.5f) * (cov[0] + cov[2]);
        const S l1 = mid + O::sqrt(O::max(O::lit(0.1f), mid * mid - det));
        const S l2 = mid - O::sqrt(O::max(O::lit(0.1f), mid * mid - det));
        const S my_radius = O::ceil(S(3) * O::sqrt(O::max(l1, l2)));
        const S pix[2] = {O::ndc2pix(S(O::d(ppx)), sc.W), O::ndc2pix(S(O::d(ppy)), sc.H)};
        int x0, y0, x1, y1;
        const int ri = (int)O::d(my_radius);
        get_rect(pix[0], pix[1], ri, gx, gy, x0, y0, x1, y1);
        if ((x1 - x0) * (y1 - y0) == 0) continue;
        if (!sc.colors_precomp) {
            S col[3]; uint8_t cl[3];
            sh_color(sc.D, sc.M, p, sc.campos, sc.shs + (size_t)i * sc.M * 3, col, cl);
            for (int c = 0; c < 3; ++c) { g.rgb[3 * (size_t)i + c] = S(O::d(col[c])); g.clamped[3 * (size_t)i + c] = cl[c]; }
        } else {
            for (int c = 0; c < 3; ++c) g.rgb[3 * (size_t)i + c] = S(sc.colors_precomp[3 * (size_t)i + c]);
        }
        g.depth[i] = S(O::d(pv[2]));
        g.radii[i] = ri;
        g.mx[i] = pix[0]; g.my[i] = pix[1];
        g.ca[i] = S(O::d(conic[0])); g.cb[i] = S(O::d(conic[1])); g.cc[i] = S(O::d(conic[2])); g.op[i] = S(sc.opacities[i]);
        g.tiles[i] = (uint32_t)((y1 - y0) * (x1 - x0));
    }
}

// duplicateWithKeys + stable SortPairs + identifyTileRanges (rasterizer_impl.cu:70-138, 304-319).
// depth bits always come from the FP32 depth (that is what the key holds).
template <class S> void bin_and_sort(const Scene& sc, const Geom<S>& g, std::vector<uint64_t>& keys, std::vector<uint32_t>& vals,
                                     std::vector<uint32_t>& ranges /*2T*/) {
    typedef Ops<S> O;
    const int gx = (sc.W + 15) / 16, gy = (sc.H + 15) / 16;
    std::vector<std::pair<uint64_t, uint32_t>> kv;
    for (int i = 0; i < sc.P; ++i) {
        if (g.radii[i] <= 0) continue;
        int x0, y0, x1, y1;
        get_rect(g.mx[i], g.my[i], g.radii[i], gx, gy, x0, y0, x1, y1);
        const float df = O::f(g.depth[i]);
        uint32_t db; std::memcpy(&db, &df, 4);
        for (int y = y0; y < y1; ++y)
            for (int x = x0; x < x1; ++x) kv.emplace_back(((uint64_t)(uint32_t)(y * gx + x) << 32) | db, (uint32_t)i);
    }
    std::stable_sort(kv.begin(), kv.end(), [](const std::pair<uint64_t, uint32_t>& a, const std::pair<uint64_t, uint32_t>& b) { return a.first < b.first; });
    const size_t R = kv.size();
    keys.resize(R); vals.resize(R);
    for (size_t i = 0; i < R; ++i) { keys[i] = kv[i].first; vals[i] = kv[i].second; }
    ranges.assign(2 * (size_t)gx * gy, 0u);
    for (size_t i = 0; i < R; ++i) {
        const uint32_t cur = (uint32_t)(keys[i] >> 32);
        if (i == 0) ranges[2 * cur] = 0;
        else {
            const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
            if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; }
        }
        if (i == R - 1) ranges[2 * cur + 1] = (uint32_t)R;
    }
}

// renderCUDA forward (forward.cu:261-379), one pixel at a time.
template <class S> void blend_forward(const Scene& sc, const Geom<S>& g, const std::vector<uint32_t>& vals,
                                      const std::vector<uint32_t>& ranges, S* out_color /*3HW*/, S* out_depth, S* final_T,
                                      uint32_t* n_contrib) {
    typedef Ops<S> O;
    const int W = sc.W, H = sc.H, gx = (W + 15) / 16;
    const size_t HW = (size_t)W * H;
    for (int py = 0; py < H; ++py)
        for (int px = 0; px < W; ++px) {
            const int tile = (py / 16) * gx + (px / 16);
            const uint32_t lo = ranges[2 * tile], hi = ranges[2 * tile + 1];
            S T = S(1), C[3] = {S(0), S(0), S(0)}, Dp = S(0);
            uint32_t contributor = 0, last = 0;
            const S pxf = S(px), pyf = S(py);
            for (uint32_t k = lo; k < hi; ++k) {
                const uint32_t id = vals[k];
                ++contributor;
                const S dx = g.mx[id] - pxf, dy = g.my[id] - pyf;
                const S power = O::lit(-0.5f) * (g.ca[id] * dx * dx + g.cc[id] * dy * dy) - g.cb[id] * dx * dy;
                if (power > S(0)) continue;
                const S alpha = O::min(O::lit(0.99f), S(O::d(g.op[id] * O::exp(S(O::d(power))))));
                if (alpha < O::lit(1.0f / 255.0f)) continue;
                const S test_T = S(O::d(T * (S(1) - alpha)));
                if (test_T < O::lit(0.0001f)) break;      // done = true
                for (int c = 0; c < 3; ++c) C[c] = C[c] + g.rgb[3 * (size_t)id + c] * alpha * T;
                Dp = Dp + g.depth[id] * alpha * T;
                T = test_T;
                last = contributor;
            }
            const size_t pix = (size_t)py * W + px;
            final_T[pix] = T;
            n_contrib[pix] = last;
            for (int c = 0; c < 3; ++c) out_color[c * HW + pix] = C[c] + T * S(sc.bg[c]);
            out_depth[pix] = Dp;
        }
}

// renderCUDA backward (backward.cu:399-557); accumulation in double for both scalar types.
template <class S> void blend_backward(const Scene& sc, const Geom<S>& g, const std::vector<uint32_t>& vals,
                                       const std::vector<uint32_t>& ranges, const S* final_T, const uint32_t* n_contrib,
                                       const float* dL_dpix, double* dmean2D /*3P*/, double* dconic /*4P*/, double* dopacity,
                                       double* dcolor /*3P*/) {
    typedef Ops<S> O;
    const int W = sc.W, H = sc.H, gx = (W + 15) / 16;
    const size_t HW = (size_t)W * H;
    const S ddelx_dx = S(0.5 * W), ddely_dy = S(0.5 * H);
    for (int py = 0; py < H; ++py)
        for (int px = 0; px < W; ++px) {
            const int tile = (py / 16) * gx + (px / 16);
            const uint32_t lo = ranges[2 * tile], hi = ranges[2 * tile + 1];
            const size_t pix = (size_t)py * W + px;
            const S T_final = final_T[pix];
            S T = T_final;
            const uint32_t last_contributor = n_contrib[pix];
            S accum_rec[3] = {S(0), S(0), S(0)}, last_color[3] = {S(0), S(0), S(0)}, last_alpha = S(0);
            const S dLp[3] = {S(dL_dpix[pix]), S(dL_dpix[HW + pix]), S(dL_dpix[2 * HW + pix])};
            S bg_dot = S(0);
            for (int c = 0; c < 3; ++c) bg_dot = bg_dot + S(sc.bg[c]) * dLp[c];
            bg_dot = S(O::d(bg_dot));
            const S pxf = S(px), pyf = S(py);
            uint32_t contributor = hi - lo;
            for (uint32_t k = hi; k > lo; --k) {
                const uint32_t id = vals[k - 1];
                --contributor;
                if (contributor >= last_contributor) continue;
                const S dx = g.mx[id] - pxf, dy = g.my[id] - pyf;
                const S power = O::lit(-0.5f) * (g.ca[id] * dx * dx + g.cc[id] * dy * dy) - g.cb[id] * dx * dy;
                if (power > S(0)) continue;
                const S G = O::exp(S(O::d(power)));
                const S alpha = O::min(O::lit(0.99f), S(O::d(g.op[id] * G)));
                if (alpha < O::lit(1.0f / 255.0f)) continue;
                T = T / (S(1) - alpha);
                const S dchannel_dcolor = S(O::d(alpha * T));
                S dL_dalpha = S(0);
                for (int c = 0; c < 3; ++c) {
                    const S col = g.rgb[3 * (size_t)id + c];
                    accum_rec[c] = S(O::d(last_alpha * last_color[c] + (S(1) - last_alpha) * accum_rec[c]));
                    last_color[c] = col;
                    dL_dalpha = dL_dalpha + (col - accum_rec[c]) * dLp[c];
                    dcolor[3 * (size_t)id + c] += O::d(dchannel_dcolor * dLp[c]);
                }
                dL_dalpha = S(O::d(dL_dalpha * T));
                last_alpha = alpha;
                dL_dalpha = S(O::d(dL_dalpha + (-T_final / (S(1) - alpha)) * bg_dot));
                const S dL_dG = S(O::d(g.op[id] * dL_dalpha));
                const S gdx = S(O::d(G * dx)), gdy = S(O::d(G * dy));
                const S dG_ddelx = S(O::d(-gdx * g.ca[id] - gdy * g.cb[id]));
                const S dG_ddely = S(O::d(-gdy * g.cc[id] - gdx * g.cb[id]));
                dmean2D[3 * (size_t)id + 0] += O::d(dL_dG * dG_ddelx * ddelx_dx);
                dmean2D[3 * (size_t)id + 1] += O::d(dL_dG * dG_ddely * ddely_dy);
                dconic[4 * (size_t)id + 0] += O::d(O::lit(-0.5f) * gdx * dx * dL_dG);
                dconic[4 * (size_t)id + 1] += O::d(O::lit(-0.5f) * gdx * dy * dL_dG);
                dconic[4 * (size_t)id + 3] += O::d(O::lit(-0.5f) * gdy * dy * dL_dG);
                dopacity[id] += O::d(G * dL_dalpha);
            }
        }
}

// computeCov2DCUDA (backward.cu:144-274) + preprocessCUDA backward (backward.cu:346-396) +
// computeColorFromSH backward (:20-139) + computeCov3D backward (:278-341).
template <class S> void preprocess_backward(const Scene& sc, const Geom<S>& g, const double* dmean2D, const double* dconic,
                                            const double* dcolor, double* dmeans3D, double* dcov3D, double* dsh,
                                            double* dscale, double* drot) {
    typedef Ops<S> O;
    const S fy = S(sc.H / (2.0f * sc.tanfovy)), fx = S(sc.W / (2.0f * sc.tanfovx));
    const float* V = sc.view; const float* Pm = sc.proj;
    for (int i = 0; i < sc.P; ++i) {
        if (!(g.radii[i] > 0)) continue;
        const S mean[3] = {S(sc.means3D[3 * i]), S(sc.means3D[3 * i + 1]), S(sc.means3D[3 * i + 2])};
        S c6[6];
        for (int k = 0; k < 6; ++k) c6[k] = sc.cov3D_precomp ? S(sc.cov3D_precomp[6 * (size_t)i + k]) : g.cov3D[6 * (size_t)i + k];
        S cov[3], t[3], xm, ym; Mat3<S> T;
        cov2d(mean, fx, fy, S(sc.tanfovx), S(sc.tanfovy), c6, V, cov, &T, t, &xm, &ym);
        const S a = cov[0], b = cov[1], c = cov[2];
        const S gA = S(dconic[4 * (size_t)i]), gB = S(dconic[4 * (size_t)i + 1]), gC = S(dconic[4 * (size_t)i + 3]);
        const S denom = a * c - b * b;
        const S denom2inv = S(1) / ((denom * denom) + O::lit(0.0000001f));
        S da = S(0), db = S(0), dc = S(0);
        const S u[3] = {T.m[0][0], T.m[0][1], T.m[0][2]}, v[3] = {T.m[1][0], T.m[1][1], T.m[1][2]};
        S dcv[6] = {S(0), S(0), S(0), S(0), S(0), S(0)};
        if (denom2inv != S(0)) {
            da = denom2inv * (-c * c * gA + S(2) * b * c * gB + (denom - a * c) * gC);
            dc = denom2inv * (-a * a * gC + S(2) * a * b * gB + (denom - a * c) * gA);
            db = denom2inv * S(2) * (b * c * gA - (denom + S(2) * b * b) * gB + a * b * gC);
            dcv[0] = u[0] * u[0] * da + u[0] * v[0] * db + v[0] * v[0] * dc;
            dcv[3] = u[1] * u[1] * da + u[1] * v[1] * db + v[1] * v[1] * dc;
            dcv[5] = u[2] * u[2] * da + u[2] * v[2] * db + v[2] * v[2] * dc;
            dcv[1] = S(2) * u[0] * u[1] * da + (u[0] * v[1] + u[1] * v[0]) * db + S(2) * v[0] * v[1] * dc;
            dcv[2] = S(2) * u[0] * u[2] * da + (u[0] * v[2] + u[2] * v[0]) * db + S(2) * v[0] * v[2] * dc;
            dcv[4] = S(2) * u[2] * u[1] * da + (u[1] * v[2] + u[2] * v[1]) * db + S(2) * v[1] * v[2] * dc;
        }
        for (int k = 0; k < 6; ++k) dcov3D[6 * (size_t)i + k] = O::d(dcv[k]);
        const S Vr[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
        S Vu[3], Vv[3];
        for (int r = 0; r < 3; ++r) {
            Vu[r] = u[0] * Vr[r][0] + u[1] * Vr[r][1] + u[2] * Vr[r][2];
            Vv[r] = v[0] * Vr[r][0] + v[1] * Vr[r][1] + v[2] * Vr[r][2];
        }
        S dT0[3], dT1[3];
        for (int r = 0; r < 3; ++r) {
            dT0[r] = S(2) * Vu[r] * da + Vv[r] * db;
            dT1[r] = S(2) * Vv[r] * dc + Vu[r] * db;
        }
        // W[0] = (V0,V4,V8), W[1] = (V1,V5,V9), W[2] = (V2,V6,V10)
        const S dJ00 = S(V[0]) * dT0[0] + S(V[4]) * dT0[1] + S(V[8]) * dT0[2];
        const S dJ02 = S(V[2]) * dT0[0] + S(V[6]) * dT0[1] + S(V[10]) * dT0[2];
        const S dJ11 = S(V[1]) * dT1[0] + S(V[5]) * dT1[1] + S(V[9]) * dT1[2];
        const S dJ12 = S(V[2]) * dT1[0] + S(V[6]) * dT1[1] + S(V[10]) * dT1[2];
        const S tz = S(1) / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
        const S dtx = xm * -fx * tz2 * dJ02;
        const S dty = ym * -fy * tz2 * dJ12;
        const S dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (S(2) * fx * t[0]) * tz3 * dJ02 + (S(2) * fy * t[1]) * tz3 * dJ12;
        S dm[3] = {S(V[0]) * dtx + S(V[1]) * dty + S(V[2]) * dtz, S(V[4]) * dtx + S(V[5]) * dty + S(V[6]) * dtz,
                   S(V[8]) * dtx + S(V[9]) * dty + S(V[10]) * dtz};
        // mean2D -> mean3D (backward.cu:366-387)
        {
            S mh[4];
            xf44(mean, Pm, mh);
            const S mw = S(1) / (mh[3] + O::lit(0.0000001f));
            const S mul1 = mh[0] * mw * mw, mul2 = mh[1] * mw * mw;
            const S g2x = S(dmean2D[3 * (size_t)i]), g2y = S(dmean2D[3 * (size_t)i + 1]);
            for (int k = 0; k < 3; ++k)
                dm[k] = dm[k] + ((S(Pm[4 * k]) * mw - S(Pm[4 * k + 3]) * mul1) * g2x + (S(Pm[4 * k + 1]) * mw - S(Pm[4 * k + 3]) * mul2) * g2y);
        }
        // SH backward (backward.cu:20-139)
        if (sc.shs) {
            const S dirO[3] = {mean[0] - S(sc.campos[0]), mean[1] - S(sc.campos[1]), mean[2] - S(sc.campos[2])};
            const S len = O::sqrt(dirO[0] * dirO[0] + dirO[1] * dirO[1] + dirO[2] * dirO[2]);
            const S x = dirO[0] / len, y = dirO[1] / len, z = dirO[2] / len;
            const float* sh = sc.shs + (size_t)i * sc.M * 3;
            double* ds = dsh + (size_t)i * sc.M * 3;
            S gcol[3];
            for (int ch = 0; ch < 3; ++ch) gcol[ch] = g.clamped[3 * (size_t)i + ch] ? S(0) : S(dcolor[3 * (size_t)i + ch]);
            S bas[16], bx[16], by[16], bz[16];
            for (int k = 0; k < 16; ++k) { bas[k] = S(0); bx[k] = S(0); by[k] = S(0); bz[k] = S(0); }
            int nact = 1;
            bas[0] = S(SH_C0);
            if (sc.D > 0) {
                nact = 4;
                bas[1] = S(-SH_C1) * y; by[1] = S(-SH_C1);
                bas[2] = S(SH_C1) * z;  bz[2] = S(SH_C1);
                bas[3] = S(-SH_C1) * x; bx[3] = S(-SH_C1);
                if (sc.D > 1) {
                    nact = 9;
                    const S xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                    bas[4] = S(SH_C2[0]) * xy; bx[4] = S(SH_C2[0]) * y; by[4] = S(SH_C2[0]) * x;
                    bas[5] = S(SH_C2[1]) * yz; by[5] = S(SH_C2[1]) * z; bz[5] = S(SH_C2[1]) * y;
                    bas[6] = S(SH_C2[2]) * (S(2) * zz - xx - yy); bx[6] = S(SH_C2[2]) * S(-2) * x; by[6] = S(SH_C2[2]) * S(-2) * y; bz[6] = S(SH_C2[2]) * S(4) * z;
                    bas[7] = S(SH_C2[3]) * xz; bx[7] = S(SH_C2[3]) * z; bz[7] = S(SH_C2[3]) * x;
                    bas[8] = S(SH_C2[4]) * (xx - yy); bx[8] = S(SH_C2[4]) * S(2) * x; by[8] = S(SH_C2[4]) * S(-2) * y;
                    if (sc.D > 2) {
                        nact = 16;
                        bas[9] = S(SH_C3[0]) * y * (S(3) * xx - yy); bx[9] = S(SH_C3[0]) * S(6) * xy; by[9] = S(SH_C3[0]) * S(3) * (xx - yy);
                        bas[10] = S(SH_C3[1]) * xy * z; bx[10] = S(SH_C3[1]) * yz; by[10] = S(SH_C3[1]) * xz; bz[10] = S(SH_C3[1]) * xy;
                        bas[11] = S(SH_C3[2]) * y * (S(4) * zz - xx - yy); bx[11] = S(SH_C3[2]) * S(-2) * xy;
                        by[11] = S(SH_C3[2]) * (S(-3) * yy + S(4) * zz - xx); bz[11] = S(SH_C3[2]) * S(8) * yz;
                        bas[12] = S(SH_C3[3]) * z * (S(2) * zz - S(3) * xx - S(3) * yy); bx[12] = S(SH_C3[3]) * S(-6) * xz;
                        by[12] = S(SH_C3[3]) * S(-6) * yz; bz[12] = S(SH_C3[3]) * S(3) * (S(2) * zz - xx - yy);
                        bas[13] = S(SH_C3[4]) * x * (S(4) * zz - xx - yy); bx[13] = S(SH_C3[4]) * (S(-3) * xx + S(4) * zz - yy);
                        by[13] = S(SH_C3[4]) * S(-2) * xy; bz[13] = S(SH_C3[4]) * S(8) * xz;
                        bas[14] = S(SH_C3[5]) * z * (xx - yy); bx[14] = S(SH_C3[5]) * S(2) * xz; by[14] = S(SH_C3[5]) * S(-2) * yz; bz[14] = S(SH_C3[5]) * (xx - yy);
                        bas[15] = S(SH_C3[6]) * x * (xx - S(3) * yy); bx[15] = S(SH_C3[6]) * S(3) * (xx - yy); by[15] = S(SH_C3[6]) * S(-6) * xy;
                    }
                }
            }
            S dd[3] = {S(0), S(0), S(0)};
            for (int k = 0; k < nact; ++k)
                for (int ch = 0; ch < 3; ++ch) {
                    ds[3 * k + ch] = O::d(bas[k] * gcol[ch]);
                    const S sg = S(sh[3 * k + ch]) * gcol[ch];
                    dd[0] = dd[0] + bx[k] * sg; dd[1] = dd[1] + by[k] * sg; dd[2] = dd[2] + bz[k] * sg;
                }
            // dnormvdv (auxiliary.h:99-110)
            const S s2 = dirO[0] * dirO[0] + dirO[1] * dirO[1] + dirO[2] * dirO[2];
            const S inv = S(1) / O::sqrt(s2 * s2 * s2);
            dm[0] = dm[0] + ((s2 - dirO[0] * dirO[0]) * dd[0] - dirO[1] * dirO[0] * dd[1] - dirO[2] * dirO[0] * dd[2]) * inv;
            dm[1] = dm[1] + (-dirO[0] * dirO[1] * dd[0] + (s2 - dirO[1] * dirO[1]) * dd[1] - dirO[2] * dirO[1] * dd[2]) * inv;
            dm[2] = dm[2] + (-dirO[0] * dirO[2] * dd[0] - dirO[1] * dirO[2] * dd[1] + (s2 - dirO[2] * dirO[2]) * dd[2]) * inv;
        }
        for (int k = 0; k < 3; ++k) dmeans3D[3 * (size_t)i + k] = O::d(dm[k]);
        // cov3D -> scale / rotation (backward.cu:278-341)
        if (sc.scales) {
            const S s3[3] = {S(sc.scales[3 * i]), S(sc.scales[3 * i + 1]), S(sc.scales[3 * i + 2])};
            const S q4[4] = {S(sc.rotations[4 * i]), S(sc.rotations[4 * i + 1]), S(sc.rotations[4 * i + 2]), S(sc.rotations[4 * i + 3])};
            S tmp[6]; Mat3<S> R, M;
            cov3d(s3, S(sc.scale_modifier), q4, tmp, &R, &M);
            const S s[3] = {S(sc.scale_modifier) * s3[0], S(sc.scale_modifier) * s3[1], S(sc.scale_modifier) * s3[2]};
            Mat3<S> dSg;
            const S h = O::lit(0.5f);
            dSg.m[0][0] = dcv[0]; dSg.m[0][1] = h * dcv[1]; dSg.m[0][2] = h * dcv[2];
            dSg.m[1][0] = h * dcv[1]; dSg.m[1][1] = dcv[3]; dSg.m[1][2] = h * dcv[4];
            dSg.m[2][0] = h * dcv[2]; dSg.m[2][1] = h * dcv[4]; dSg.m[2][2] = dcv[5];
            Mat3<S> M2;
            for (int cI = 0; cI < 3; ++cI) for (int rI = 0; rI < 3; ++rI) M2.m[cI][rI] = S(2) * M.m[cI][rI];
            Mat3<S> dM = mul(M2, dSg);
            S n[3][3];
            for (int a2 = 0; a2 < 3; ++a2) {
                dscale[3 * (size_t)i + a2] = O::d(R.m[0][a2] * dM.m[0][a2] + R.m[1][a2] * dM.m[1][a2] + R.m[2][a2] * dM.m[2][a2]);
                for (int b2 = 0; b2 < 3; ++b2) n[a2][b2] = dM.m[b2][a2] * s[a2];
            }
            const S r = q4[0], x = q4[1], y = q4[2], z = q4[3];
            drot[4 * (size_t)i + 0] = O::d(S(2) * z * (n[0][1] - n[1][0]) + S(2) * y * (n[2][0] - n[0][2]) + S(2) * x * (n[1][2] - n[2][1]));
            drot[4 * (size_t)i + 1] = O::d(S(2) * y * (n[1][0] + n[0][1]) + S(2) * z * (n[2][0] + n[0][2]) + S(2) * r * (n[1][2] - n[2][1]) - S(4) * x * (n[2][2] + n[1][1]));
            drot[4 * (size_t)i + 2] = O::d(S(2) * x * (n[1][0] + n[0][1]) + S(2) * r * (n[2][0] - n[0][2]) + S(2) * z * (n[1][2] + n[2][1]) - S(4) * y * (n[2][2] + n[0][0]));
            drot[4 * (size_t)i + 3] = O::d(S(2) * r * (n[0][1] - n[1][0]) + S(2) * x * (n[2][0] + n[0][2]) + S(2) * y * (n[1][2] + n[2][1]) - S(4) * z * (n[1][1] + n[0][0]));
        }
    }
}

// ---- flat C interface (ctypes) ----------------------------------------------------------------
struct Out {
    int32_t* radii; uint32_t* tiles_touched; float* depths; float* means2D; float* cov3D; float* conic_opacity; float* rgb;
    uint8_t* clamped;
    uint64_t* keys; uint32_t* point_list; uint32_t* ranges; int64_t cap;     // keys/point_list capacity
    double* color; double* depth; double* final_T; uint32_t* n_contrib;
    // backward (optional)
    const float* dL_dpix;
    double *dmean2D, *dconic, *dopacity, *dcolor, *dmeans3D, *dcov3D, *dsh, *dscale, *drot;
};

template <class S> int64_t run(const Scene& sc, Out& o) {
    typedef Ops<S> O;
    Geom<S> g;
    preprocess(sc, g);
    std::vector<uint64_t> keys; std::vector<uint32_t> vals, ranges;
    bin_and_sort(sc, g, keys, vals, ranges);
    const int64_t R = (int64_t)keys.size();
    const size_t P = sc.P, HW = (size_t)sc.W * sc.H;
    for (size_t i = 0; i < P; ++i) {
        o.radii[i] = g.radii[i]; o.tiles_touched[i] = g.tiles[i]; o.depths[i] = O::f(g.depth[i]);
        o.means2D[2 * i] = O::f(g.mx[i]); o.means2D[2 * i + 1] = O::f(g.my[i]);
        for (int k = 0; k < 6; ++k) o.cov3D[6 * i + k] = O::f(g.cov3D[6 * i + k]);
        o.conic_opacity[4 * i] = O::f(g.ca[i]); o.conic_opacity[4 * i + 1] = O::f(g.cb[i]);
        o.conic_opacity[4 * i + 2] = O::f(g.cc[i]); o.conic_opacity[4 * i + 3] = O::f(g.op[i]);
        for (int k = 0; k < 3; ++k) { o.rgb[3 * i + k] = O::f(g.rgb[3 * i + k]); o.clamped[3 * i + k] = g.clamped[3 * i + k]; }
    }
    for (size_t i = 0; i < ranges.size(); ++i) o.ranges[i] = ranges[i];
    if (R > o.cap) return -R;
    for (int64_t i = 0; i < R; ++i) { o.keys[i] = keys[i]; o.point_list[i] = vals[i]; }
    std::vector<S> col(3 * HW), dep(HW), fT(HW);
    blend_forward(sc, g, vals, ranges, col.data(), dep.data(), fT.data(), o.n_contrib);
    for (size_t i = 0; i < 3 * HW; ++i) o.color[i] = O::d(col[i]);
    for (size_t i = 0; i < HW; ++i) { o.depth[i] = O::d(dep[i]); o.final_T[i] = O::d(fT[i]); }
    if (o.dL_dpix) {
        blend_backward(sc, g, vals, ranges, fT.data(), o.n_contrib, o.dL_dpix, o.dmean2D, o.dconic, o.dopacity, o.dcolor);
        preprocess_backward(sc, g, o.dmean2D, o.dconic, o.dcolor, o.dmeans3D, o.dcov3D, o.dsh, o.dscale, o.drot);
    }
    return R;
}

}  // namespace

extern "C" int64_t fdgs_oracle_raster(int use_f64, int P, int D, int M, int W, int H, float tanfovx, float tanfovy,
                                      float scale_modifier, const float* bg, const float* means3D, const float* shs,
                                      const float* colors_precomp, const float* opacities, const float* scales,
                                      const float* rotations, const float* cov3D_precomp, const float* view,
                                      const float* proj, const float* campos, Out* out) {
    Scene sc{P, D, M, W, H, tanfovx, tanfovy, scale_modifier, bg, means3D, shs, colors_precomp, opacities, scales, rotations,
             cov3D_precomp, view, proj, campos};
    return use_f64 ? run<double>(sc, *out) : run<Ff>(sc, *out);
}
