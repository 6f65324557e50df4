// rt_imgdiff.cpp — image comparison for the tolerance tiers of SURVEY.md §8(c) and the
// "display / read-back" row §8(f3): RMSE and mean LDR-FLIP between two tone-mapped images.
//
// The reference has no image output and no comparison tool (its texture is only ever consumed
// by shaders/render.wgsl:34-39), so there is nothing upstream to mirror: this file restates the
// published LDR-FLIP metric (Andersson et al., "FLIP: A Difference Evaluator for Alternating
// Images", HPG 2020) from the paper's equations: contrast-sensitivity filtering in YCxCz,
// Hunt-adjusted CIELab + HyAB colour difference with the paper's error redistribution
// (qc = 0.7, pc = 0.4, pt = 0.95), edge / point feature differences on the achromatic
// channel (w = 0.082, qf = 0.5), FLIP = dEc^(1 - dEf).  NVIDIA's own implementation is not in
// this image, so the numbers are pinned only by an independent numpy/scipy restatement
// (tests/test_imgdiff.py) and by closed-form cases (identical images, black vs white).
//
// Inputs are what rt_read_rgba8 / Context::save_png produce: sRGB-encoded values in [0, 1],
// 3 floats per pixel, any row order (both images the same).  Arithmetic in double.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <stdexcept>
#include <vector>
#include "rt_host.hpp"

namespace rtb {
namespace {

const double kPi = 3.14159265358979323846;

struct C3 { double a, b, c; };

double srgb_to_linear(double v) { return v > 0.04045 ? std::pow((v + 0.055) / 1.055, 2.4) : v / 12.92; }

C3 linrgb_to_xyz(C3 p)
{
    return {(10135552.0 * p.a + 8788810.0 * p.b + 4435075.0 * p.c) / 24577794.0,
            (2613072.0 * p.a + 8788810.0 * p.b + 887015.0 * p.c) / 12288897.0,
            (1425312.0 * p.a + 8788810.0 * p.b + 70074185.0 * p.c) / 73733382.0};
}
C3 xyz_to_linrgb(C3 p)
{
    return {3.241003275 * p.a - 1.537398934 * p.b - 0.498615861 * p.c,
            -0.969224334 * p.a + 1.875930071 * p.b + 0.041554224 * p.c,
            0.055639423 * p.a - 0.204011202 * p.b + 1.057148933 * p.c};
}
C3 illuminant() { return linrgb_to_xyz({1.0, 1.0, 1.0}); }
C3 xyz_to_ycxcz(C3 p)
{
    const C3 w = illuminant();
    const double x = p.a / w.a, y = p.b / w.b, z = p.c / w.c;
    return {116.0 * y - 16.0, 500.0 * (x - y), 200.0 * (y - z)};
}
C3 ycxcz_to_xyz(C3 p)
{
    const C3 w = illuminant();
    const double y = (p.a + 16.0) / 116.0, x = y + p.b / 500.0, z = y - p.c / 200.0;
    return {x * w.a, y * w.b, z * w.c};
}
double lab_f(double t)
{
    const double delta = 6.0 / 29.0;
    return t > 0.00885 ? std::cbrt(t) : t / (3.0 * delta * delta) + 4.0 / 29.0;
}
C3 xyz_to_lab(C3 p)
{
    const C3 w = illuminant();
    const double x = lab_f(p.a / w.a), y = lab_f(p.b / w.b), z = lab_f(p.c / w.c);
    return {116.0 * y - 16.0, 500.0 * (x - y), 200.0 * (y - z)};
}
C3 hunt(C3 lab) { return {lab.a, 0.01 * lab.a * lab.b, 0.01 * lab.a * lab.c}; }
double hyab(C3 p, C3 q) { return std::fabs(p.a - q.a) + std::sqrt((p.b - q.b) * (p.b - q.b) + (p.c - q.c) * (p.c - q.c)); }

// (2r+1)^2 kernel, replicate border
void convolve(const std::vector<double>& src, std::vector<double>& dst, int w, int h, const std::vector<double>& k, int r)
{
    dst.assign((size_t)w * h, 0.0);
#pragma omp parallel for schedule(static)
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            double s = 0.0;
            for (int dy = -r; dy <= r; dy++) {
                const int yy = std::min(std::max(y + dy, 0), h - 1);
                const double* row = &src[(size_t)yy * w];
                const double* kr = &k[(size_t)(dy + r) * (2 * r + 1)];
                for (int dx = -r; dx <= r; dx++) s += kr[dx + r] * row[std::min(std::max(x + dx, 0), w - 1)];
            }
            dst[(size_t)y * w + x] = s;
        }
}

std::vector<double> csf_kernel(double a1, double b1, double a2, double b2, int r, double ppd)
{
    std::vector<double> k((size_t)(2 * r + 1) * (2 * r + 1));
    double sum = 0.0;
    for (int y = -r; y <= r; y++)
        for (int x = -r; x <= r; x++) {
            const double d = (x / ppd) * (x / ppd) + (y / ppd) * (y / ppd);
            const double g = a1 * std::sqrt(kPi / b1) * std::exp(-kPi * kPi * d / b1) + a2 * std::sqrt(kPi / b2) * std::exp(-kPi * kPi * d / b2);
            k[(size_t)(y + r) * (2 * r + 1) + (x + r)] = g; sum += g;
        }
    for (double& v : k) v /= sum;
    return k;
}

// x-derivative detectors: edge = first, point = second derivative of a Gaussian; positive and
// negative weights normalised to +1 / -1 separately.  The y detector is the transpose.
std::vector<double> feature_kernel(bool point, int r, double sd, bool transpose)
{
    const int n = 2 * r + 1;
    std::vector<double> k((size_t)n * n);
    double pos = 0.0, neg = 0.0;
    for (int y = -r; y <= r; y++)
        for (int x = -r; x <= r; x++) {
            const double g = std::exp(-(double)(x * x + y * y) / (2.0 * sd * sd));
            const double v = point ? ((double)(x * x) / (sd * sd) - 1.0) * g : -(double)x * g;
            k[(size_t)(y + r) * n + (x + r)] = v;
            if (v > 0.0) pos += v; else neg -= v;
        }
    for (double& v : k) v = v > 0.0 ? v / pos : (v < 0.0 ? v / neg : 0.0);
    if (transpose) {
        std::vector<double> t((size_t)n * n);
        for (int y = 0; y < n; y++) for (int x = 0; x < n; x++) t[(size_t)x * n + y] = k[(size_t)y * n + x];
        return t;
    }
    return k;
}

struct Planes { std::vector<double> p[3]; };

Planes to_ycxcz(const float* srgb, size_t n)
{
    Planes o; for (auto& v : o.p) v.resize(n);
#pragma omp parallel for schedule(static)
    for (long long i = 0; i < (long long)n; i++) {
        const C3 lin = {srgb_to_linear(std::min(std::max((double)srgb[3 * i], 0.0), 1.0)),
                        srgb_to_linear(std::min(std::max((double)srgb[3 * i + 1], 0.0), 1.0)),
                        srgb_to_linear(std::min(std::max((double)srgb[3 * i + 2], 0.0), 1.0))};
        const C3 c = xyz_to_ycxcz(linrgb_to_xyz(lin));
        o.p[0][i] = c.a; o.p[1][i] = c.b; o.p[2][i] = c.c;
    }
    return o;
}

} // namespace

double image_rmse(const float* a, const float* b, size_t n_values)
{
    if (n_values == 0) return 0.0;
    double s = 0.0;
    for (size_t i = 0; i < n_values; i++) { const double d = (double)a[i] - (double)b[i]; s += d * d; }
    return std::sqrt(s / (double)n_values);
}

double image_flip(const float* ref_srgb, const float* test_srgb, uint32_t width, uint32_t height, double ppd, float* error_map)
{
    if (width == 0 || height == 0) throw std::runtime_error("image_flip: empty image");
    if (!(ppd > 0.0)) throw std::runtime_error("image_flip: pixels per degree must be positive");
    const int w = (int)width, h = (int)height;
    const size_t n = (size_t)w * h;
    const Planes R = to_ycxcz(ref_srgb, n), T = to_ycxcz(test_srgb, n);

    // ---- colour pipeline: CSF filtering, clamp in linear RGB, Hunt-adjusted Lab, HyAB ----
    const int rc = (int)std::ceil(3.0 * std::sqrt(0.04 / (2.0 * kPi * kPi)) * ppd);
    const std::vector<double> kA = csf_kernel(1.0, 0.0047, 0.0, 1e-5, rc, ppd), kRG = csf_kernel(1.0, 0.0053, 0.0, 1e-5, rc, ppd),
                              kBY = csf_kernel(34.1, 0.04, 13.5, 0.025, rc, ppd);
    Planes FR, FT;
    convolve(R.p[0], FR.p[0], w, h, kA, rc); convolve(R.p[1], FR.p[1], w, h, kRG, rc); convolve(R.p[2], FR.p[2], w, h, kBY, rc);
    convolve(T.p[0], FT.p[0], w, h, kA, rc); convolve(T.p[1], FT.p[1], w, h, kRG, rc); convolve(T.p[2], FT.p[2], w, h, kBY, rc);
    const double qc = 0.7, pc = 0.4, pt = 0.95;
    const double cmax = std::pow(hyab(hunt(xyz_to_lab(linrgb_to_xyz({0.0, 1.0, 0.0}))), hunt(xyz_to_lab(linrgb_to_xyz({0.0, 0.0, 1.0})))), qc);
    const double pccmax = pc * cmax;
    std::vector<double> dEc(n);
    auto perceptual = [](C3 ycc) {
        C3 lin = xyz_to_linrgb(ycxcz_to_xyz(ycc));
        lin = {std::min(std::max(lin.a, 0.0), 1.0), std::min(std::max(lin.b, 0.0), 1.0), std::min(std::max(lin.c, 0.0), 1.0)};
        return hunt(xyz_to_lab(linrgb_to_xyz(lin)));
    };
#pragma omp parallel for schedule(static)
    for (long long i = 0; i < (long long)n; i++) {
        const double d = std::pow(hyab(perceptual({FR.p[0][i], FR.p[1][i], FR.p[2][i]}), perceptual({FT.p[0][i], FT.p[1][i], FT.p[2][i]})), qc);
        dEc[i] = d < pccmax ? (pt / pccmax) * d : pt + ((d - pccmax) / (cmax - pccmax)) * (1.0 - pt);
    }

    // ---- feature pipeline on the unfiltered achromatic channel ----
    const double sd = 0.5 * 0.082 * ppd, qf = 0.5;
    const int rf = (int)std::ceil(3.0 * sd);
    std::vector<double> yr(n), yt(n);
    for (size_t i = 0; i < n; i++) { yr[i] = (R.p[0][i] + 16.0) / 116.0; yt[i] = (T.p[0][i] + 16.0) / 116.0; }
    std::vector<double> mag[2][2];                 // [edge/point][ref/test]
    for (int f = 0; f < 2; f++)
        for (int s = 0; s < 2; s++) {
            std::vector<double> gx, gy;
            convolve(s ? yt : yr, gx, w, h, feature_kernel(f == 1, rf, sd, false), rf);
            convolve(s ? yt : yr, gy, w, h, feature_kernel(f == 1, rf, sd, true), rf);
            mag[f][s].resize(n);
            for (size_t i = 0; i < n; i++) mag[f][s][i] = std::sqrt(gx[i] * gx[i] + gy[i] * gy[i]);
        }
    double sum = 0.0;
    for (size_t i = 0; i < n; i++) {
        const double de = std::fabs(mag[0][0][i] - mag[0][1][i]), dp = std::fabs(mag[1][0][i] - mag[1][1][i]);
        const double dEf = std::pow(std::max(de, dp) / std::sqrt(2.0), qf);
        const double v = std::pow(dEc[i], 1.0 - dEf);
        if (error_map) error_map[i] = (float)v;
        sum += v;
    }
    return sum / (double)n;
}

} // namespace rtb
