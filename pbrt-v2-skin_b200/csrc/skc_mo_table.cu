// skc_mo_table.cu -- host-side builder and verifier of the compact Rd table (skc_mo_table.h).
#include <algorithm>
#include <cstring>
#include <atomic>
#include <thread>
#include "skc_mo_table.h"
#include "../../include/skin_cuda.h"

static double cubicAt(const MoCompactTable &C, int l, float d2) {
    int k; float w[4];
    mocCoord(C.P, d2, k, w);
    if (k < 0 || k > C.P.K - 2) return NAN;
    const float *r = C.rows.data() + (size_t)k * 32 + l;
    return (double)mocCubic(w, r[0], r[32], r[64], r[96]);
}

bool mocBuild(const float *table, const float *rcp, uint32_t len, float tol, int K, int M, uint32_t headCapFloats, MoCompactTable &C) {
    if (len < 16 || K < 8 || K > MOC_ROWS_MAX || M < 1) return false;
    float rcpMin = INFINITY;
    for (int l = 0; l < SKC_NSPECTRAL; ++l) rcpMin = std::min(rcpMin, rcp[l]);
    // abscissa range: the smooth part of the longest table ends well before its last third (fp32 noise of the
    // dipole sums, see the header); u = K-1 sits at 0.3 of its d2 range
    const double xhi = 0.3 * (double)(len - 1) / (double)rcpMin;
    const double binades = (double)(K - 1) / (double)M;
    const double a = xhi / (exp2(binades) - 1.0);
    C.P.invA = (float)(1.0 / a); C.P.cellsPerBinade = (float)M; C.P.K = K;
    C.tol = tol;
    C.rows.assign((size_t)(K + 3) * 32, 0.f);
    const double aEff = 1.0 / (double)C.P.invA;
    std::vector<double> xs(K + 3);
    for (int u = -1; u <= K + 1; ++u) xs[u + 1] = aEff * (exp2((double)u / (double)M) - 1.0);
    bool any = false;
    uint32_t headUsed = 0;
    double covSum = 0.0;
    for (int l = 0; l < 32; ++l) { C.headOff[l] = 0; C.headLen[l] = 0; C.fLo[l] = 0.f; C.fHi[l] = 0.f; C.maxErr[l] = 0.f; }
    // the wavelengths are independent (each owns one column of C.rows and its own window): rows and verification on worker threads,
    // then the shared-memory heads are handed out in wavelength order
    bool haveWindow[32] = {false};
    auto oneWavelength = [&](int l) {
        std::vector<double> v(K + 3);
        const float *data = table + (size_t)l * len;
        const double r = (double)rcp[l];
        // the last table segment may fall to a 0 end marker: sample the rows from the segments before it
        const double fEnd = (double)(len - 2);
        int lo = -1, hi = -1;
        for (int i = 0; i < K + 3; ++i) {
            double f = xs[i] * r;
            if (xs[i] >= 0.0 && f <= fEnd) {
                uint32_t seg = (uint32_t)f; if (seg > len - 3) seg = len - 3;
                double t = f - (double)seg;
                v[i] = (1.0 - t) * (double)data[seg] + t * (double)data[seg + 1];
                if (lo < 0) lo = i;
                hi = i;
            }
        }
        if (lo < 0 || hi - lo < 4) return;
        // rows outside the table's own range continue the cubic through the last four inside it
        for (int i = lo - 1; i >= 0; --i) v[i] = 4 * v[i + 1] - 6 * v[i + 2] + 4 * v[i + 3] - v[i + 4];
        for (int i = hi + 1; i < K + 3; ++i) v[i] = 4 * v[i - 1] - 6 * v[i - 2] + 4 * v[i - 3] - v[i - 4];
        for (int i = 0; i < K + 3; ++i) C.rows[(size_t)i * 32 + l] = (float)v[i];
        // verify: four points per table segment, the device's own arithmetic on an fp32 d2
        const float fCap = (float)std::min(fEnd, xs[K] * r * 0.999);      // abscissa K-1 is the last with rows on both sides
        static const double offs[4] = {0.0, 0.27, 0.5, 0.77};
        double bestLo = 0, bestHi = -1, runLo = -1, runMax = 0, bestMax = 0;
        double lastGood = -1;
        auto closeRun = [&]() {
            if (runLo >= 0 && lastGood - runLo > bestHi - bestLo) { bestLo = runLo; bestHi = lastGood; bestMax = runMax; }
            runLo = -1; runMax = 0;
        };
        for (uint32_t i = 0; i + 2 < len; ++i) {
            bool stop = false;
            for (int o = 0; o < 4; ++o) {
                float d2 = (float)(((double)i + offs[o]) / r);
                float f = d2 * rcp[l];
                if (f >= fCap) { stop = true; break; }
                double ex = (double)mocExact(data, len, rcp[l], d2);
                double cu = cubicAt(C, l, d2);
                double err = ex != 0.0 ? fabs(cu - ex) / fabs(ex) : (cu == 0.0 ? 0.0 : 1.0);
                if (err == err && err <= 0.9 * tol) {            // 0.9: margin for MUFU.LG2 against log2f and for the points in between
                    if (runLo < 0) runLo = f;
                    lastGood = f; runMax = std::max(runMax, err);
                } else closeRun();
            }
            if (stop) break;
        }
        closeRun();
        if (bestHi - bestLo < 64.0) return;                       // not worth a window
        // one segment of margin either side
        float fLo = bestLo <= 0.0 ? 0.f : (float)ceil(bestLo + 1.0), fHi = (float)floor(bestHi - 1.0);
        C.fLo[l] = fLo; C.fHi[l] = fHi; C.maxErr[l] = (float)bestMax;
        haveWindow[l] = true;
    };
    {
        std::atomic<int> next(0);
        auto work = [&]() { for (int l = next.fetch_add(1); l < SKC_NSPECTRAL; l = next.fetch_add(1)) oneWavelength(l); };
        unsigned nThreads = std::min(std::max(std::thread::hardware_concurrency(), 1u), (unsigned)SKC_NSPECTRAL);
        std::vector<std::thread> pool;
        for (unsigned t = 1; t < nThreads; ++t) pool.emplace_back(work);
        work();
        for (std::thread &t : pool) t.join();
    }
    for (int l = 0; l < SKC_NSPECTRAL; ++l) {
        if (!haveWindow[l]) continue;
        const float *data = table + (size_t)l * len;
        const float fLo = C.fLo[l], fHi = C.fHi[l];
        any = true;
        // head of the exact table below the window, in shared memory as far as the budget goes
        uint32_t want = (uint32_t)fLo;
        if (want) {
            uint32_t room = headCapFloats > headUsed + 1 ? headCapFloats - headUsed - 1 : 0;
            uint32_t hl = std::min(want, room);
            if (hl >= 32) {
                C.headOff[l] = headUsed; C.headLen[l] = hl;
                C.heads.insert(C.heads.end(), data, data + hl + 1);
                headUsed += hl + 1;
            }
        }
        covSum += (log(std::max((double)fHi, 3.0) / std::max((double)fLo, 3.0)) + log(std::max((double)C.headLen[l], 3.0) / 3.0)) / log((double)len / 3.0);
    }
    C.coverage = covSum / SKC_NSPECTRAL;
    if (C.heads.empty()) C.heads.push_back(0.f);
    return any;
}

float mocEvalHost(const MoCompactTable &C, const float *table, const float *rcp, uint32_t len, int l, float d2) {
    float f = d2 * rcp[l];
    if (f >= (float)(len - 1)) return 0.f;
    if (f >= C.fLo[l] && f < C.fHi[l]) return (float)cubicAt(C, l, d2);
    return mocExact(table + (size_t)l * len, len, rcp[l], d2);     // head or tail: the same exact rule
}

// ---- host-only test hooks (no GPU needed): tests/test_mo_table.py proves the bound over all (wavelength, d2) ------
extern "C" {
struct skc_moc_handle { MoCompactTable C; std::vector<float> table, rcp; uint32_t len; };

void *skc_moc_build(const float *table, const float *rcp, uint32_t len, float tol, int K, int M, uint32_t headCapFloats) {
    skc_moc_handle *h = new skc_moc_handle();
    h->table.assign(table, table + (size_t)SKC_NSPECTRAL * len); h->rcp.assign(rcp, rcp + SKC_NSPECTRAL); h->len = len;
    if (!mocBuild(h->table.data(), h->rcp.data(), len, tol, K, M, headCapFloats, h->C)) { delete h; return nullptr; }
    return h;
}
void skc_moc_free(void *hv) { delete (skc_moc_handle *)hv; }
// info[l] = {fLo, fHi, headLen, maxErr} per wavelength; returns coverage
double skc_moc_info(void *hv, float *info4x30, uint32_t *rowsOut, uint32_t *headFloatsOut) {
    skc_moc_handle *h = (skc_moc_handle *)hv;
    for (int l = 0; l < SKC_NSPECTRAL; ++l) {
        info4x30[4 * l] = h->C.fLo[l]; info4x30[4 * l + 1] = h->C.fHi[l]; info4x30[4 * l + 2] = (float)h->C.headLen[l]; info4x30[4 * l + 3] = h->C.maxErr[l];
    }
    if (rowsOut) *rowsOut = (uint32_t)h->C.P.K + 3;
    if (headFloatsOut) *headFloatsOut = (uint32_t)h->C.heads.size();
    return h->C.coverage;
}
// out[i] = what the kernel computes for wavelength l at d2[i]; exact[i] = sampleProfile
void skc_moc_eval(void *hv, int l, const float *d2, uint64_t n, float *out, float *exact) {
    skc_moc_handle *h = (skc_moc_handle *)hv;
    for (uint64_t i = 0; i < n; ++i) {
        out[i] = mocEvalHost(h->C, h->table.data(), h->rcp.data(), h->len, l, d2[i]);
        if (exact) exact[i] = mocExact(h->table.data() + (size_t)l * h->len, h->len, h->rcp[l], d2[i]);
    }
}
}
