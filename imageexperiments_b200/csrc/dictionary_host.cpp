// Dictionary builder, host only: what the dictionary half of compressed::createCompressionContextFast
// computes at start-up (CompressedImage.cpp:317-341) - the distinct line-cut shapes of a block (BasisSet.cpp:204-297), the
// blurred +-1 segment atoms (createSegmentDictionaryFast, :430-511) and, per shape and colour channel, the KLT refinement rows
// of the two segments from a parametric covariance model (createIntraSegmentDictionaryFast, :618-721; createBasisFast,
// :154-186; models :12-24).
//
// Numerics. Shapes are integer geometry and the segment atoms are float arithmetic in a fixed order: both are reproduced
// exactly (tests compare them bit for bit with the dictionary exported from the reference build). The reference then runs
// Eigen's single-precision SelfAdjointEigenSolver on near-rank-1 covariance matrices, whose rounding depends on the Eigen
// version and build - that part cannot be matched bit for bit and is PINNED here instead: cyclic Jacobi in
// double on the float covariance matrix, fixed sweep order, eigenvectors rounded to float once at the end. Streams written with
// this dictionary form their own codec revision: they decode with this dictionary, not with a reference build's.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "../../include/mnenc.h"
#include "mnenc_internal.h"

namespace mnenc {

int fail(int code, const std::string& msg);  // mnenc_api.cu
int abiCaught() noexcept;

namespace {

struct Pt {
    int x, y;
};
struct Line {
    Pt a, b;
};

// sign test of the reference's Dist(line.a, line.b, p) >= 0 (BasisSet.cpp:188-190): the denominator is a positive length, so the
// sign is that of the integer cross product; a zero-length line gives 0 / 0 = NaN there, and NaN >= 0 is false.
bool onPositiveSide(const Line& l, Pt p) {
    if (l.a.x == l.b.x && l.a.y == l.b.y) return false;
    const long cross = static_cast<long>(l.b.x - l.a.x) * (l.a.y - p.y) - static_cast<long>(l.a.x - p.x) * (l.b.y - l.a.y);
    return cross >= 0;
}

using Bits = std::vector<bool>;
struct ShapeOrder {  // at the first differing pixel the shape with the pixel SET comes first (ShapeComparator, :192-202)
    bool operator()(const Bits& a, const Bits& b) const {
        for (size_t i = 0; i < a.size(); ++i)
            if (a[i] != b[i]) return a[i];
        return false;
    }
};

Bits shapeOf(const Line& l, int bs, bool invert) {
    Bits bits(static_cast<size_t>(bs) * bs);
    for (int x = 0; x < bs; ++x)
        for (int y = 0; y < bs; ++y) bits[x + y * bs] = onPositiveSide(l, Pt{x, y}) != invert;
    return bits;
}

// The distinct ways a straight line cuts a bs x bs block, one line per shape, in shape order (510 for bs = 8).
std::vector<Line> lineShapes(int bs) {
    std::map<Bits, Line, ShapeOrder> shapes;
    auto offer = [&](const Line& l, bool alsoInverse) {
        const Bits bits = shapeOf(l, bs, false);
        if (shapes.count(bits)) return;
        if (alsoInverse && shapes.count(shapeOf(l, bs, true))) return;
        shapes.emplace(bits, l);
    };
    // axis-parallel cuts first (and with them the two constant shapes): sloped lines that cut the same pixels come later and lose
    for (int side = -1; side < bs + 1; ++side) {
        offer(Line{Pt{0, side}, Pt{bs, side}}, false);
        offer(Line{Pt{side, 0}, Pt{side, bs}}, false);
    }
    // lines from the row below the block to the left column / the row above / the right column of the 3 bs x 3 bs surround
    for (int s1 = -bs; s1 < 2 * bs; ++s1)
        for (int s2 = -bs; s2 < 2 * bs; ++s2) {
            offer(Line{Pt{s1, -bs}, Pt{-bs, s2}}, true);
            offer(Line{Pt{s1, -bs}, Pt{s2, bs}}, true);
            offer(Line{Pt{s1, -bs}, Pt{bs, s2}}, true);
        }
    // lines from the right column to the left column / the row above (the reference offers the second family twice, :271-291)
    for (int s1 = -bs; s1 < 2 * bs; ++s1)
        for (int s2 = -bs; s2 < 2 * bs; ++s2) {
            offer(Line{Pt{bs, s1}, Pt{-bs, s2}}, true);
            offer(Line{Pt{bs, s1}, Pt{s2, bs}}, true);
        }
    std::vector<Line> out;
    out.reserve(shapes.size());
    for (const auto& kv : shapes) out.push_back(kv.second);
    return out;
}

// One segment atom: the +-1 mask of the line on a doubled grid, blurred by a 7 x 7 Gaussian (sigma 1) sampled at the even
// positions, mean removed unless the mask is constant, unit norm. Float arithmetic in the reference's order (:430-511).
void segmentAtom(const Line& l, int bs, float* row) {
    const int half = 3, kw = 7, zs = 2 * bs;
    float gauss[7][7];
    for (int dx = -half; dx <= half; ++dx)
        for (int dy = -half; dy <= half; ++dy) gauss[dx + half][dy + half] = static_cast<float>(std::exp(-static_cast<double>(dx * dx + dy * dy) / (2.0 * 1.0 * 1.0)));
    (void)kw;
    const Line doubled{Pt{l.a.x * 2, l.a.y * 2}, Pt{l.b.x * 2, l.b.y * 2}};
    std::vector<float> zoom(static_cast<size_t>(zs) * zs);
    bool allSet = true, allClear = true;
    for (int x = 0; x < zs; ++x)
        for (int y = 0; y < zs; ++y) {
            const bool bit = onPositiveSide(doubled, Pt{x, y});
            (bit ? allClear : allSet) = false;
            zoom[x + zs * y] = bit ? 1.0f : -1.0f;
        }
    float total = 0.0f;
    for (int x = 0; x < bs; ++x)
        for (int y = 0; y < bs; ++y) {
            float weight = 0.0f, tot = 0.0f;
            for (int dx = -half; dx <= half; ++dx) {
                const int u = std::clamp(2 * x + dx, 0, zs - 1);
                for (int dy = -half; dy <= half; ++dy) {
                    const int v = std::clamp(2 * y + dy, 0, zs - 1);
                    const float coeff = gauss[dx + half][dy + half];
                    weight += coeff;
                    tot += coeff * zoom[u + zs * v];
                }
            }
            const float value = tot / weight;
            total += value;
            row[x + bs * y] = value;
        }
    const float mean = total / static_cast<float>(bs * bs);
    float sumsq = 0.0f;
    for (int j = 0; j < bs * bs; ++j) {
        float value = row[j];
        if (!(allSet || allClear)) {
            value -= mean;
            row[j] = value;
        }
        sumsq += value * value;
    }
    const float norm = sqrtf(sumsq);
    if (sumsq != 0.0f)
        for (int j = 0; j < bs * bs; ++j) row[j] = row[j] / norm;
}

// covariance of two pixels dx, dy apart (BasisSet.cpp:12-24: the fitted models of the three channels)
double covarianceModel(int channel, double dx, double dy) {
    switch (channel) {
        case 0:
            return 3817.7299999999996 * std::exp(-1.48854e-05 * dx * dx + -1.7273e-05 * dy * dy) +
                   657.8100000000001 * std::exp(-0.0436057 * std::fabs(dx) + -0.050844400000000005 * std::fabs(dy));
        case 1:
            return 241.49 * std::exp(-0.00134755 * std::fabs(dx) + -0.00147572 * std::fabs(dy));
        default:
            return 371.87199999999996 * std::exp(-0.00147084 * std::fabs(dx) + -0.0015265799999999998 * std::fabs(dy));
    }
}

// Eigenvectors of a symmetric n x n matrix (row-major, double), cyclic Jacobi: rows of `vec` on return, eigenvalues in `val`.
void jacobiEigen(std::vector<double>& a, int n, std::vector<double>& vec, std::vector<double>& val) {
    vec.assign(static_cast<size_t>(n) * n, 0.0);
    for (int i = 0; i < n; ++i) vec[static_cast<size_t>(i) * n + i] = 1.0;
    double scale = 0.0;
    for (double x : a) scale += x * x;
    for (int sweep = 0; sweep < 64; ++sweep) {
        double off = 0.0;
        for (int p = 0; p < n; ++p)
            for (int q = p + 1; q < n; ++q) off += a[static_cast<size_t>(p) * n + q] * a[static_cast<size_t>(p) * n + q];
        if (off <= 1e-30 * scale) break;
        for (int p = 0; p < n; ++p)
            for (int q = p + 1; q < n; ++q) {
                const double apq = a[static_cast<size_t>(p) * n + q];
                if (apq == 0.0) continue;
                const double app = a[static_cast<size_t>(p) * n + p], aqq = a[static_cast<size_t>(q) * n + q];
                const double theta = (aqq - app) / (2.0 * apq);
                const double t = (theta >= 0.0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < n; ++k) {  // columns p, q
                    const double akp = a[static_cast<size_t>(k) * n + p], akq = a[static_cast<size_t>(k) * n + q];
                    a[static_cast<size_t>(k) * n + p] = c * akp - s * akq;
                    a[static_cast<size_t>(k) * n + q] = s * akp + c * akq;
                }
                for (int k = 0; k < n; ++k) {  // rows p, q
                    const double apk = a[static_cast<size_t>(p) * n + k], aqk = a[static_cast<size_t>(q) * n + k];
                    a[static_cast<size_t>(p) * n + k] = c * apk - s * aqk;
                    a[static_cast<size_t>(q) * n + k] = s * apk + c * aqk;
                }
                for (int k = 0; k < n; ++k) {  // eigenvectors, kept as rows
                    const double vp = vec[static_cast<size_t>(p) * n + k], vq = vec[static_cast<size_t>(q) * n + k];
                    vec[static_cast<size_t>(p) * n + k] = c * vp - s * vq;
                    vec[static_cast<size_t>(q) * n + k] = s * vp + c * vq;
                }
            }
    }
    val.resize(n);
    for (int i = 0; i < n; ++i) val[i] = a[static_cast<size_t>(i) * n + i];
}

// KLT basis of one segment (createBasisFast, :154-186): eigenvectors of the covariance of its pixels by falling |eigenvalue|,
// sign fixed so that the first significant entry is positive; float.
std::vector<std::vector<float>> segmentBasis(const std::vector<Pt>& px, int channel) {
    const int n = static_cast<int>(px.size());
    std::vector<std::vector<float>> out;
    if (n == 0) return out;
    std::vector<double> cov(static_cast<size_t>(n) * n), vec, val;
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j)  // the reference stores the model's value as float (:637-655)
            cov[static_cast<size_t>(i) * n + j] =
                static_cast<double>(static_cast<float>(covarianceModel(channel, static_cast<double>(px[i].x - px[j].x), static_cast<double>(px[i].y - px[j].y))));
    jacobiEigen(cov, n, vec, val);
    std::vector<int> order(n);
    for (int i = 0; i < n; ++i) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return std::fabs(val[x]) > std::fabs(val[y]); });
    for (int i : order) {
        std::vector<float> v(n);
        double sign = 1.0;
        for (int j = 0; j < n; ++j)
            if (std::fabs(vec[static_cast<size_t>(i) * n + j]) > 1e-10) {
                if (vec[static_cast<size_t>(i) * n + j] < 0.0) sign = -1.0;
                break;
            }
        for (int j = 0; j < n; ++j) v[j] = static_cast<float>(sign * vec[static_cast<size_t>(i) * n + j]);
        out.push_back(std::move(v));
    }
    return out;
}

// Refinement rows of one shape and channel (:618-721): for each of the two segments its KLT vectors but the first, placed on the
// segment's pixels, mean over the segment removed, unit norm; the two segments' rows interleaved. Returns the row count.
int detailRowsOf(const Line& l, int bs, int channel, float* rows) {
    std::vector<Pt> seg[2];
    for (int x = 0; x < bs; ++x)
        for (int y = 0; y < bs; ++y) seg[onPositiveSide(l, Pt{x, y}) ? 0 : 1].push_back(Pt{x, y});
    const std::vector<std::vector<float>> basis[2] = {segmentBasis(seg[0], channel), segmentBasis(seg[1], channel)};
    const int count[2] = {std::max(0, static_cast<int>(seg[0].size()) - 1), std::max(0, static_cast<int>(seg[1].size()) - 1)};
    const int cols = bs * bs;
    int offset = 0;
    for (int i = 0; i < std::max(count[0], count[1]); ++i)
        for (int s = 0; s < 2; ++s) {
            if (i >= count[s]) continue;
            float* row = rows + static_cast<size_t>(offset) * cols;
            std::fill(row, row + cols, 0.0f);
            const std::vector<float>& v = basis[s][i + 1];
            float meanTotal = 0.0f;
            for (size_t j = 0; j < seg[s].size(); ++j) {
                meanTotal += v[j];
                row[seg[s][j].x + bs * seg[s][j].y] = v[j];
            }
            meanTotal /= static_cast<float>(seg[s].size());
            float sumSq = 0.0f;
            for (size_t j = 0; j < seg[s].size(); ++j) {
                float& value = row[seg[s][j].x + bs * seg[s][j].y];
                value -= meanTotal;
                sumSq += value * value;
            }
            sumSq = std::sqrt(sumSq);
            if (s == 0 || sumSq != 0.0f)  // the reference guards the division for the second segment only (:703-707)
                for (int j = 0; j < cols; ++j) row[j] /= sumSq;
            ++offset;
        }
    return offset;
}

}  // namespace

}  // namespace mnenc

extern "C" int mnenc_build_dictionary(int block_size, int threads, float* base_rowmajor, int* n_base, int32_t* detail_rows, float* detailY,
                                      float* detailU, float* detailV) try {
    using namespace mnenc;
    if (!base_rowmajor || !n_base || !detail_rows || !detailY || !detailU || !detailV) return fail(MNENC_ERR_ARG, "null argument");
    if (block_size != 8) return fail(MNENC_ERR_ARG, "only block_size 8 is supported");
    const int bs = block_size, cols = bs * bs;
    const std::vector<Line> shapes = lineShapes(bs);
    const int nb = static_cast<int>(shapes.size());
    if (nb > kRowsPad) return fail(MNENC_ERR_INTERNAL, "more shapes than the dictionary layout holds");
    *n_base = nb;
    for (int a = 0; a < nb; ++a) segmentAtom(shapes[a], bs, base_rowmajor + static_cast<size_t>(a) * cols);
    // every (shape, channel) is independent: 3 * 510 pairs of eigenproblems over the host threads, into slots of 64 rows each
    std::vector<float> slots(static_cast<size_t>(3) * nb * cols * cols);
    std::vector<int> rows(static_cast<size_t>(3) * nb, 0);
    int nt = threads > 0 ? threads : static_cast<int>(std::thread::hardware_concurrency());
    nt = std::clamp(nt, 1, 64);
    std::vector<std::thread> pool;
    for (int t = 0; t < nt; ++t)
        pool.emplace_back([&, t] {
            for (int item = t; item < 3 * nb; item += nt)
                rows[item] = detailRowsOf(shapes[item % nb], bs, item / nb, slots.data() + static_cast<size_t>(item) * cols * cols);
        });
    for (auto& th : pool) th.join();
    float* out[3] = {detailY, detailU, detailV};
    for (int ch = 0; ch < 3; ++ch) {
        size_t at = 0;
        for (int a = 0; a < nb; ++a) {
            const int r = rows[static_cast<size_t>(ch) * nb + a];
            if (ch == 0) detail_rows[a] = r;
            else if (detail_rows[a] != r) return fail(MNENC_ERR_INTERNAL, "channels disagree about a group's height");
            std::memcpy(out[ch] + at * cols, slots.data() + (static_cast<size_t>(ch) * nb + a) * cols * cols, static_cast<size_t>(r) * cols * sizeof(float));
            at += static_cast<size_t>(r);
        }
    }
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}
