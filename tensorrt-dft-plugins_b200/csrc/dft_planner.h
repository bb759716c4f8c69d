// dft_planner.h -- host-side planner: radix schedules, twiddle / permutation tables and the list
// of kernel launches for one (direction, rank, dims, batch).  Header-only; used by the CUDA
// executor (dft_kernels.cu) and by the test-only host emulator (tests/emu/dft_emu.cpp).
//
// This is the replacement for what cufftXtMakePlanMany does for the reference
// (/root/reference/src/dft_plugins/dft_plugins.cpp:171-176).
#pragma once
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "dft_types.h"

namespace dft {

constexpr int kSmemMax = 227 * 1024 - 64;   // usable dynamic shared memory per CTA (bytes)
constexpr int kBarrierBytes = 16;           // mbarrier slot appended after the tile
constexpr int kThreads = 256;

enum BufSel : int { BUF_IN = 0, BUF_OUT = 1, BUF_WS = 2, BUF_WS1 = 3 };

struct Launch {
    KParams kp;            // pointers left null; resolved by the executor from the ids below
    int src, dst;          // BufSel
    int tw_row_id = -1, tw_col_id = -1, tw_unt_id = -1, tw_row_t_id = -1;
    int perm_row_id = -1, perm_col_id = -1;
    int tw_four_id = -1;   // four-step first pass: w_Ntot^j table
    std::string name;
};

struct PlanDesc {
    int direction = 0, rank = 0;
    int64_t dims[3] = {1, 1, 1};
    int64_t batch = 0;
    int64_t in_floats = 0, out_floats = 0;
    size_t ws_bytes = 0;
    size_t ws1_offset = 0;   // start of the second workspace half (BUF_WS + 1), naive inverse only
    std::vector<Launch> launches;
    std::vector<std::vector<float2>> tw_tables;
    std::vector<std::vector<int>> perm_tables;
    std::string description;
};

// ---- radix schedule --------------------------------------------------------------------------
static const int kRadices[] = {16, 13, 11, 10, 9, 8, 7, 6, 5, 4, 3, 2, 17, 19, 23, 29, 31};

inline void factor_search(int n, int start, std::vector<int>& cur, std::vector<int>& best)
{
    if (n == 1) {
        bool better = best.empty() || cur.size() < best.size();
        if (!better && cur.size() == best.size()) {
            const int mc = *std::max_element(cur.begin(), cur.end());
            const int mb = *std::max_element(best.begin(), best.end());
            better = mc < mb;
        }
        if (better) best = cur;
        return;
    }
    if (!best.empty() && cur.size() + 1 > best.size()) return;
    for (int i = start; i < (int)(sizeof(kRadices) / sizeof(int)); ++i) {
        const int r = kRadices[i];
        if (n % r) continue;
        cur.push_back(r);
        factor_search(n / r, i, cur, best);
        cur.pop_back();
    }
}

inline bool is_pow2(int n) { return n > 0 && (n & (n - 1)) == 0; }
inline int phys_index(int p, bool pad) { return pad ? p + (p >> 4) + (p >> 8) : p; }

constexpr int kMaxGenericPrime = 8191;      // prime factors above 31 run as generic O(R) -per-output steps (fft_step_generic)

inline bool radix_is_generic(int r) { return r > 31; }
inline bool line_has_generic(const LineFFT& f)
{
    for (int i = 0; i < f.nsteps; ++i)
        if (radix_is_generic(f.radix[i])) return true;
    return false;
}

// false when n has a prime factor above kMaxGenericPrime or needs too many steps.
inline bool make_line_fft(int n, bool rows, LineFFT* f)
{
    memset(f, 0, sizeof(*f));
    f->n = n;
    if (n < 1) return false;
    if (n == 1) return true;
    int m = n;
    for (int p : {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31})
        while (m % p == 0) m /= p;
    // what is left are prime factors above 31: each becomes a radix of its own (a generic step)
    std::vector<int> big;
    for (int p = 37; (long long)p * p <= m; p += 2)
        while (m % p == 0) {
            big.push_back(p);
            m /= p;
        }
    if (m > 1) big.push_back(m);
    for (int p : big)
        if (p > kMaxGenericPrime) return false;
    const int small = n / [&] { long long q = 1; for (int p : big) q *= p; return (int)q; }();
    std::vector<int> cur, best;
    if (small > 1) factor_search(small, 0, cur, best);
    if ((small > 1 && best.empty()) || (int)(best.size() + big.size()) > kMaxSteps) return false;
    std::sort(best.begin(), best.end(), [](int a, int b) { return a > b; });
    if (!big.empty()) {
        // generic steps first (largest block length: their one-output-per-item form does not care), butterflies after
        std::sort(big.begin(), big.end(), [](int a, int b) { return a > b; });
        best.insert(best.begin(), big.begin(), big.end());
        f->nsteps = (int)best.size();
        int L = n;
        for (int i = 0; i < f->nsteps; ++i) {
            f->radix[i] = best[i];
            const unsigned mm = (unsigned)(L / best[i]), bpl = (unsigned)(n / best[i]);
            f->mmag[i] = mm <= 1 ? 0u : (unsigned)(((1ull << 32) + mm - 1) / mm);
            f->bmag[i] = bpl <= 1 ? 0u : (unsigned)(((1ull << 32) + bpl - 1) / bpl);
            L /= best[i];
        }
        return true;
    }
    if (rows && is_pow2(n) && n >= 32) {
        // power-of-two rows are stored skewed (phys(p) = p + p/16 + p/256); the skew is linear inside a
        // butterfly when the schedule is r_top, 16, 16, ... (n = r_top * 16^k)
        best.clear();
        int rest = n;
        while (rest > 16) {
            best.insert(best.begin(), 16);
            rest /= 16;
        }
        if (rest > 1) best.insert(best.begin(), rest);
    } else if (rows) {
        // (1) the last row step reads r contiguous elements per thread: an odd radix makes the lane stride
        //     odd in 8-byte units, i.e. shared-memory bank-conflict free;
        // (2) results are gathered in natural order k = c0 + r0*(...) from position c0*(n/r0) + ...: with all
        //     factors of two in r0, n/r0 is odd and consecutive k hit different banks.
        int pick = -1;
        for (int i = 0; i < (int)best.size(); ++i)
            if (best[i] % 2 == 1) { pick = i; break; }
        if (pick >= 0) {
            const int r = best[pick];
            best.erase(best.begin() + pick);
            best.push_back(r);
        }
        // try to put every factor of two into the first radix without adding steps
        int two = 1;
        while (n % (two * 2) == 0 && two < 16) two *= 2;
        if (two >= 2 && n % (two * 2) != 0) {
            std::vector<int> pick_list;
            for (int r0 : {16, 10, 8, 6, 4, 2}) {
                if (n % r0 || r0 % two || (r0 / two) % 2 == 0) continue;
                std::vector<int> c2, b2;
                if (n / r0 > 1) factor_search(n / r0, 0, c2, b2);
                if (n / r0 > 1 && b2.empty()) continue;
                if (b2.size() + 1 <= best.size() && (pick_list.empty() || b2.size() + 1 < pick_list.size())) {
                    std::sort(b2.begin(), b2.end(), [](int a, int b) { return a > b; });
                    pick_list = b2;
                    pick_list.insert(pick_list.begin(), r0);
                }
            }
            if (!pick_list.empty()) best = pick_list;
        }
    }
    f->nsteps = (int)best.size();
    int L = n;
    for (int i = 0; i < f->nsteps; ++i) {
        f->radix[i] = best[i];
        const unsigned m = (unsigned)(L / best[i]), bpl = (unsigned)(n / best[i]);
        f->mmag[i] = m <= 1 ? 0u : (unsigned)(((1ull << 32) + m - 1) / m);
        f->bmag[i] = bpl <= 1 ? 0u : (unsigned)(((1ull << 32) + bpl - 1) / bpl);
        L /= best[i];
    }
    return true;
}

// position (in the in-place DIF result) of frequency k
inline int dif_position(const LineFFT& f, int k)
{
    int L = f.n, pos = 0;
    for (int s = 0; s < f.nsteps; ++s) {
        const int r = f.radix[s];
        const int m = L / r;
        pos += (k % r) * m;
        k /= r;
        L = m;
    }
    return pos;
}

inline std::vector<int> make_perm(const LineFFT& f, bool inverse, bool pad = false)
{
    std::vector<int> p(f.n);
    for (int i = 0; i < f.n; ++i) p[i] = phys_index(dif_position(f, inverse ? (f.n - i) % f.n : i), pad);
    return p;
}

// w_n^j = exp(-2 pi i j / n), j in [0, count)
inline std::vector<float2> make_twiddles(int n, int count)
{
    std::vector<float2> t(count);
    for (int j = 0; j < count; ++j) {
        const double a = 2.0 * 3.14159265358979323846264338327950288 * (double)(j % n) / (double)n;
        t[j] = float2{(float)cos(a), (float)(-sin(a))};
    }
    return t;
}

// per-step transposed twiddles of a row schedule (KParams::tw_row_t)
inline std::vector<float2> make_twiddles_steps(const LineFFT& f)
{
    std::vector<float2> t;
    int L = f.n;
    for (int s = 0; s < f.nsteps; ++s) {
        const int r = f.radix[s], m = L / r;
        for (int c = 1; c < r; ++c)
            for (int b = 0; b < m; ++b) {
                const double a = 2.0 * 3.14159265358979323846264338327950288 * (double)(((long long)b * c) % L) / (double)L;
                t.push_back(float2{(float)cos(a), (float)(-sin(a))});
            }
        L = m;
    }
    if (t.empty()) t.push_back(float2{1.f, 0.f});
    return t;
}

// ---- plan -------------------------------------------------------------------------------------
enum PlanFlags : int { PLAN_NO_SMALL_ROWS = 1 };   // keep rows on the tile engine (needed for column-split output)

struct Planner {
    PlanDesc* pd;
    int flags = 0;
    int add_tw(std::vector<float2> t) { pd->tw_tables.push_back(std::move(t)); return (int)pd->tw_tables.size() - 1; }
    int add_perm(std::vector<int> t) { pd->perm_tables.push_back(std::move(t)); return (int)pd->perm_tables.size() - 1; }

    static KParams base(int mode)
    {
        KParams k;
        memset(&k, 0, sizeof(k));
        k.mode = mode;
        k.threads = kThreads;
        k.scale = 1.0f;
        k.G = 1;
        return k;
    }

    static int pick_group(long long nunits, long long unit_bytes, long long target_bytes)
    {
        long long g = target_bytes / std::max<long long>(unit_bytes, 1);
        const long long spread = (nunits + 148 * 4 - 1) / (148 * 4);   // keep >= ~4 tiles per SM
        g = std::min(g, spread);
        g = std::min<long long>(g, (long long)kSmemMax / std::max<long long>(unit_bytes, 1));
        return (int)std::max<long long>(g, 1);
    }

    // rows pass (1-D R2C / C2R over `nrows` rows of real length W)
    void add_rows(bool inverse, long long nrows, int W, int src, int dst, float scale)
    {
        const bool even = (W % 2 == 0);
        const int M = even ? W / 2 : W;
        // A handful of short rows is launch-latency bound: one thread per row, all in registers (BASELINE config 1).
        if (even && (M == 2 || M == 4 || M == 8 || M == 16 || M == 32) && nrows <= 8192 && !(flags & PLAN_NO_SMALL_ROWS) && !getenv("TRT_DFT_NO_SMALL_ROWS")) {
            KParams k = base(inverse ? MODE_SMALL_ROWS_C2R : MODE_SMALL_ROWS_R2C);
            k.W = W;
            k.M = M;
            k.H = 1;
            k.nunits = nrows;
            k.ntiles = (nrows + 63) / 64;
            k.threads = 64;
            k.smem_bytes = 0;
            k.scale = scale;
            Launch l;
            l.kp = k;
            l.src = src;
            l.dst = dst;
            char nm[96];
            snprintf(nm, sizeof nm, "%s<W=%d> rows=%lld [one thread per row, registers only]", inverse ? "small_rows_c2r" : "small_rows_r2c", W, nrows);
            l.name = nm;
            pd->launches.push_back(l);
            return;
        }
        KParams k = base(inverse ? (even ? MODE_I1D : MODE_I1D_ODD) : (even ? MODE_F1D : MODE_F1D_ODD));
        k.W = W;
        k.M = M;
        k.H = 1;
        make_line_fft(M, true, &k.row);
        k.pad = (even && is_pow2(M) && M >= 32) ? 1 : 0;
        k.P = k.pad ? phys_index(M, true) + 1 : ((inverse && even) ? M + 1 : M);
        if (even && !k.pad && !line_has_generic(k.row) && !getenv("TRT_DFT_NO_ROW_ORDER")) set_row_orders(k.row, k.P);   // bank-conflict-free work-item order
        k.nunits = nrows;
        const bool gen = line_has_generic(k.row);      // generic prime steps need a second copy of the tile
        k.G = pick_group(nrows, (long long)k.P * 8 * (gen ? 2 : 1), 48 * 1024);
        k.ntiles = (nrows + k.G - 1) / k.G;
        k.smem_bytes = k.G * k.P * 8 * (gen ? 2 : 1) + kBarrierBytes;
        k.scratch_off = gen ? k.G * k.P : 0;
        k.scale = scale;
        Launch l;
        l.kp = k;
        l.src = src;
        l.dst = dst;
        l.tw_row_id = add_tw(make_twiddles(M, M));
        l.tw_row_t_id = add_tw(make_twiddles_steps(k.row));
        l.perm_row_id = add_perm(make_perm(k.row, inverse, k.pad != 0));
        if (even) l.tw_unt_id = add_tw(make_twiddles(W, M + 1));
        char nm[128];
        snprintf(nm, sizeof nm, "%s<W=%d> rows=%lld G=%d smem=%d", inverse ? (even ? "rows_c2r" : "rows_c2r_odd") : (even ? "rows_r2c" : "rows_r2c_odd"),
                 W, nrows, k.G, k.smem_bytes);
        l.name = nm;
        pd->launches.push_back(l);
    }

    // Long strided axes (a column tile of >= 4 x 8 bytes no longer fits twice in shared memory) run as a
    // four-step FFT, N = N1*N2: pass 1 gathers rows n1 + N1*n2, transforms over n2 (length N2), multiplies
    // by w_N^(n1*k2) and stores them as the contiguous rows n1*N2 + k2 of the workspace; pass 2 transforms
    // over n1 (length N1, row stride N2) and leaves natural order.  Both passes use full-width coalesced
    // tiles.  The inverse reads pass 1's rows index-reversed instead of using reversed output tables.
    bool add_cols_four_step(bool inverse, long long outer, int N, long long inner, int src, int dst, float scale)
    {
        if (src == BUF_WS || getenv("TRT_DFT_NO_FOURSTEP")) return false;
        if ((long long)N * 8 * 4 * 2 <= kSmemMax - 1024) return false;     // a 4-wide tile still fits twice: one pass is fine
        int best1 = 0;
        for (int n2 = 2; n2 <= 2048; ++n2) {
            if (N % n2) continue;
            const int n1 = N / n2;
            LineFFT f1, f2;
            if (n1 < 2 || n1 > 2048 || !make_line_fft(n1, false, &f1) || !make_line_fft(n2, false, &f2)) continue;
            auto bal = [](int a, int b) { return a > b ? (double)a / b : (double)b / a; };
            if (!best1 || bal(n1, n2) < bal(best1, N / best1) || (bal(n1, n2) == bal(best1, N / best1) && n2 > N / best1)) best1 = n1;
        }
        if (!best1) return false;
        const int N1 = best1, N2 = N / best1;
        // The workspace rows are padded to an even number of complex elements: a 16-byte multiple, which makes the
        // workspace a legal 2-D TMA tensor for pass 2 (the native one-sided pitch, (W/2+1)*8 bytes, never is).
        // (not when the second pass runs in place on the workspace -- the un-fused inverse -- where it would read the padded
        // layout and write the plain one over it)
        const bool pad_ok = dst != BUF_WS && inner <= (1 << 28) && !getenv("TRT_DFT_NO_WS_PAD");
        const long long inner_p = pad_ok ? inner + (inner & 1) : inner;
        const size_t bytes = (size_t)outer * N * inner_p * 8;
        pd->ws_bytes = std::max(pd->ws_bytes, bytes);
        const size_t before = pd->launches.size();
        add_cols(false, outer * N1, N2, inner, src, BUF_WS, 1.0f, /*allow_four_step=*/false);
        {
            Launch& l = pd->launches[before];
            l.kp.n1split = N1;
            l.kp.rev_in = inverse ? 1 : 0;
            l.kp.out_rowlen = pad_ok ? (int)inner_p : 0;
            l.tw_four_id = add_tw(make_twiddles(N, N));
            l.name += " [four-step pass 1 of " + std::to_string(N) + "]";
        }
        add_cols(false, outer, N1, (long long)N2 * inner_p, BUF_WS, dst, scale, false);
        {
            Launch& l = pd->launches.back();
            if (pad_ok) {
                l.kp.unflat_in = (int)inner_p;
                l.kp.unflat_w = (int)inner;
                // one tensor copy per tile: box = [G columns][N1 rows], both at most 256
                l.kp.tma_cols = (N1 <= 256 && l.kp.G <= 128 && l.kp.G == l.kp.P && !getenv("TRT_DFT_NO_TMA_COLS")) ? 1 : 0;
            }
            l.name += pad_ok ? " [four-step pass 2, padded workspace]" : " [four-step pass 2]";
        }
        return true;
    }

    // complex FFT of length N along a strided axis of data[outer][N][inner]
    void add_cols(bool inverse, long long outer, int N, long long inner, int src, int dst, float scale, bool allow_four_step = true)
    {
        if (N == 1 && src == dst && scale == 1.0f) return;   // identity
        if (allow_four_step && add_cols_four_step(inverse, outer, N, inner, src, dst, scale)) return;
        KParams k = base(MODE_C2C_COLS);
        k.H = N;
        make_line_fft(N, false, &k.col);
        const bool gen = line_has_generic(k.col);
        const long long maxfit = std::max<long long>(1, (long long)(kSmemMax - kBarrierBytes) / ((long long)N * 8 * (gen ? 2 : 1)));
        long long tm = std::max(std::min<long long>(16, maxfit), std::min<long long>(32, (64 * 1024) / ((long long)N * 8)));
        // prefer a tile that leaves room for two pipeline stages, as long as rows stay >= 32 bytes wide
        {
            long long t2 = 16;
            while (t2 > 1 && t2 * N * 8 * (gen ? 2 : 1) > (kSmemMax - 1024) / 2) t2 /= 2;
            // short axes: widen the tile to as much as 64 KB -- per-tile overheads (barriers, pipeline handshakes) amortise
            // and the row pieces fetched from DRAM get longer (128-point first pass of cfg5's four-step columns: 256-byte
            // pieces at a 2 MB stride with 32 columns, 512-byte pieces with 64: 1.45 -> 1.42 ms forward, 1.72 -> 1.63 inverse)
            while (t2 < 64 && t2 * 2 * N * 8 <= 64 * 1024) t2 *= 2;
            if (t2 >= 4) tm = t2;
        }
        if (const char* e = getenv("TRT_DFT_TM")) tm = atoll(e);          // tuning override (benchmarks only)
        tm = std::max<long long>(1, std::min(tm, std::min(inner, maxfit)));
        k.G = (int)tm;
        k.P = (int)tm;
        k.inner = inner;
        k.outer = outer;
        k.chunks = (inner + tm - 1) / tm;
        k.nunits = outer * k.chunks;
        k.ntiles = k.nunits;
        k.smem_bytes = N * k.P * 8 * (gen ? 2 : 1) + kBarrierBytes;
        k.scratch_off = gen ? N * k.P : 0;
        k.scale = scale;
        k.inverse = inverse ? 1 : 0;
        k.threads = ((long long)N * tm >= 8192) ? 512 : 256;   // big tiles: more warps to hide the strided loads
        if (const char* e = getenv("TRT_DFT_COLS_THREADS")) k.threads = atoi(e) == 512 ? 512 : 256;
        Launch l;
        l.kp = k;
        l.src = src;
        l.dst = dst;
        l.tw_col_id = add_tw(make_twiddles(N, N));
        l.perm_col_id = add_perm(make_perm(k.col, inverse));
        char nm[128];
        snprintf(nm, sizeof nm, "cols_c2c<N=%d> outer=%lld inner=%lld TM=%d smem=%d", N, outer, inner, k.G, k.smem_bytes);
        l.name = nm;
        pd->launches.push_back(l);
    }

    void add_fused2d(bool inverse, long long nimg, int H, int W, int src, int dst, float scale)
    {
        const int M = W / 2;
        KParams k = base(inverse ? MODE_I2D : MODE_F2D);
        k.H = H;
        k.W = W;
        k.M = M;
        k.P = inverse ? M + 1 : M;
        make_line_fft(M, true, &k.row);
        make_line_fft(H, false, &k.col);
        const bool gen = line_has_generic(k.row) || line_has_generic(k.col);
        const long long unit = (long long)H * k.P * 8 * (gen ? 2 : 1);
        k.nunits = nimg;
        k.G = pick_group(nimg, unit, 48 * 1024);
        k.ntiles = (nimg + k.G - 1) / k.G;
        k.smem_bytes = (int)(k.G * unit) + kBarrierBytes;
        k.scratch_off = gen ? k.G * H * k.P : 0;
        k.scale = scale;
        Launch l;
        l.kp = k;
        l.src = src;
        l.dst = dst;
        l.tw_row_id = add_tw(make_twiddles(M, M));
        l.tw_row_t_id = add_tw(make_twiddles_steps(k.row));
        l.tw_col_id = add_tw(make_twiddles(H, H));
        l.tw_unt_id = add_tw(make_twiddles(W, M + 1));
        l.perm_row_id = add_perm(make_perm(k.row, inverse));
        l.perm_col_id = add_perm(make_perm(k.col, inverse));
        char nm[128];
        snprintf(nm, sizeof nm, "%s<%d,%d> images=%lld G=%d smem=%d", inverse ? "fused_irfft2" : "fused_rfft2", H, W, nimg, k.G, k.smem_bytes);
        l.name = nm;
        pd->launches.push_back(l);
    }

    void add_naive(int mode, long long total, int W, int N, long long inner, int tw_n, bool inverse, int src, int dst, float scale)
    {
        KParams k = base(mode);
        k.W = W;
        k.H = N;
        k.inner = inner;
        k.nunits = total;
        k.ntiles = total;
        k.scale = scale;
        Launch l;
        l.kp = k;
        l.src = src;
        l.dst = dst;
        if (mode == MODE_NAIVE_COLS) {
            l.kp.W = inverse ? 1 : 0;
            l.tw_col_id = add_tw(make_twiddles(tw_n, tw_n));
        } else {
            l.tw_row_id = add_tw(make_twiddles(tw_n, tw_n));
        }
        char nm[96];
        snprintf(nm, sizeof nm, "naive_mode%d<n=%d> elems=%lld", mode, tw_n, total);
        l.name = nm;
        pd->launches.push_back(l);
    }
};

inline bool line_supported(int n, bool rows_pitch_plus1)
{
    LineFFT f;
    if (!make_line_fft(n, true, &f)) return false;
    const long long pitch = (rows_pitch_plus1 && is_pow2(n) && n >= 32) ? phys_index(n, true) + 1 : n + (rows_pitch_plus1 ? 1 : 0);
    return pitch * 8 * (line_has_generic(f) ? 2 : 1) + kBarrierBytes <= kSmemMax;
}
inline bool length_has_generic(int n)
{
    LineFFT f;
    return make_line_fft(n, false, &f) && line_has_generic(f);
}

// Returns 0 on success, nonzero TRT_DFT_ERR_* style code (1 = invalid argument).
inline int build_plan(PlanDesc* pd, int direction, int rank, const int64_t* dims, int64_t batch, int flags = 0)
{
    if (!pd || !dims || rank < 1 || rank > 3 || (direction != 0 && direction != 1) || batch < 0) return 1;
    for (int i = 0; i < rank; ++i)
        if (dims[i] < 1 || dims[i] > (1LL << 30)) return 1;
    pd->direction = direction;
    pd->rank = rank;
    pd->batch = batch;
    for (int i = 0; i < 3; ++i) pd->dims[i] = i < rank ? dims[i] : 1;
    const bool inv = direction == 1;
    const int W = (int)dims[rank - 1];
    const int Mo = W / 2 + 1;
    const int H = rank >= 2 ? (int)dims[rank - 2] : 1;
    const int D = rank == 3 ? (int)dims[0] : 1;
    const long long lead = (long long)H * D;
    const long long real_floats = batch * lead * W;
    const long long cplx_floats = batch * lead * Mo * 2;
    pd->in_floats = inv ? cplx_floats : real_floats;
    pd->out_floats = inv ? real_floats : cplx_floats;
    pd->ws_bytes = 0;
    if (batch == 0) {
        pd->description = "empty";
        return 0;
    }
    double total = (double)W;
    if (rank >= 2) total *= H;
    if (rank == 3) total *= D;
    const float inv_scale = (float)(1.0 / total);
    const bool even = (W % 2 == 0);
    const int M = even ? W / 2 : W;

    Planner P{pd, flags};
    bool engine = line_supported(M, true);
    if (rank >= 2) engine = engine && line_supported(H, false);
    if (rank == 3) engine = engine && line_supported(D, false);

    if (!engine) {
        // any length: naive O(N^2) passes, ping-ponging through workspace for the strided axes
        const size_t cbytes = (size_t)cplx_floats * 4;
        if (!inv) {
            P.add_naive(MODE_NAIVE_ROWS_R2C, batch * lead * Mo, W, 1, 1, W, false, BUF_IN, BUF_OUT, 1.f);
            if (rank >= 2 && H > 1) {
                pd->ws_bytes = cbytes;
                // out -> ws (axis H), then ws -> out (axis D, or an identity pass of length 1)
                P.add_naive(MODE_NAIVE_COLS, batch * lead * Mo, 0, H, Mo, H, false, BUF_OUT, BUF_WS, 1.f);
                P.add_naive(MODE_NAIVE_COLS, batch * lead * Mo, 0, D, (long long)H * Mo, D, false, BUF_WS, BUF_OUT, 1.f);
            } else if (rank == 3 && D > 1) {
                pd->ws_bytes = cbytes;
                P.add_naive(MODE_NAIVE_COLS, batch * lead * Mo, 0, D, (long long)H * Mo, D, false, BUF_OUT, BUF_WS, 1.f);
                P.add_naive(MODE_NAIVE_COLS, batch * lead * Mo, 0, 1, 1, 1, false, BUF_WS, BUF_OUT, 1.f);
            }
        } else {
            if (rank == 1 || lead == 1) {
                P.add_naive(MODE_NAIVE_ROWS_C2R, batch * lead * W, W, 1, 1, W, true, BUF_IN, BUF_OUT, inv_scale);
            } else {
                // in -> ws[0] (axis D) -> ws[1] (axis H) -> out (rows); ws holds two complex tensors
                pd->ws1_offset = (cbytes + 255) / 256 * 256;
                pd->ws_bytes = pd->ws1_offset + cbytes;
                P.add_naive(MODE_NAIVE_COLS, batch * lead * Mo, 0, D, (long long)H * Mo, D, true, BUF_IN, BUF_WS, 1.f);
                P.add_naive(MODE_NAIVE_COLS, batch * lead * Mo, 0, H, Mo, H, true, BUF_WS, BUF_WS1, 1.f);
                P.add_naive(MODE_NAIVE_ROWS_C2R, batch * lead * W, W, 1, 1, W, true, BUF_WS1, BUF_OUT, inv_scale);
            }
        }
    } else {
        const int gen2 = (rank >= 2 && (length_has_generic(M) || length_has_generic(H))) ? 2 : 1;   // generic prime steps: tile + scratch
        const bool fuse_fwd = rank >= 2 && even && ((long long)H * M * 8 * gen2 + kBarrierBytes <= kSmemMax);
        const bool fuse_inv = rank >= 2 && even && ((long long)H * (M + 1) * 8 * gen2 + kBarrierBytes <= kSmemMax);
        if (!inv) {
            if (rank == 1) {
                P.add_rows(false, batch, W, BUF_IN, BUF_OUT, 1.f);
            } else {
                if (fuse_fwd) {
                    P.add_fused2d(false, batch * D, H, W, BUF_IN, BUF_OUT, 1.f);
                } else {
                    P.add_rows(false, batch * lead, W, BUF_IN, BUF_OUT, 1.f);
                    P.add_cols(false, batch * D, H, Mo, BUF_OUT, BUF_OUT, 1.f);
                }
                if (rank == 3) P.add_cols(false, batch, D, (long long)H * Mo, BUF_OUT, BUF_OUT, 1.f);
            }
        } else {
            if (rank == 1) {
                P.add_rows(true, batch, W, BUF_IN, BUF_OUT, inv_scale);
            } else {
                int cur = BUF_IN;
                if (rank == 3 && D > 1) {
                    pd->ws_bytes = (size_t)cplx_floats * 4;
                    P.add_cols(true, batch, D, (long long)H * Mo, cur, BUF_WS, 1.f);
                    cur = BUF_WS;
                }
                if (fuse_inv) {
                    P.add_fused2d(true, batch * D, H, W, cur, BUF_OUT, inv_scale);
                } else {
                    if (H > 1) {
                        pd->ws_bytes = (size_t)cplx_floats * 4;
                        P.add_cols(true, batch * D, H, Mo, cur, BUF_WS, 1.f);
                        cur = BUF_WS;
                    }
                    P.add_rows(true, batch * lead, W, cur, BUF_OUT, inv_scale);
                }
            }
        }
    }
    std::string d;
    for (size_t i = 0; i < pd->launches.size(); ++i) {
        if (i) d += " ; ";
        d += pd->launches[i].name;
    }
    pd->description = d;
    return 0;
}

// Complex-to-complex FFT of length n along the middle axis of data[outer][n][inner] (unnormalised;
// `scale` is applied when the result is stored).  This is the strided-axis pass of a multi-dimensional
// transform on its own, which the pencil-decomposed multi-GPU rfft2 runs after its all-to-all.
inline int build_c2c_plan(PlanDesc* pd, int direction, int64_t n, int64_t outer, int64_t inner, float scale)
{
    if (!pd || (direction != 0 && direction != 1) || n < 1 || n > (1LL << 30) || outer < 0 || inner < 1) return 1;
    pd->direction = direction;
    pd->rank = 1;
    pd->dims[0] = n;
    pd->batch = outer * inner;
    pd->in_floats = pd->out_floats = outer * n * inner * 2;
    pd->ws_bytes = 0;
    if (outer == 0) {
        pd->description = "empty";
        return 0;
    }
    Planner P{pd};
    LineFFT f;
    const bool engine = make_line_fft((int)n, false, &f) && (long long)n * 8 * (line_has_generic(f) ? 2 : 1) + kBarrierBytes <= kSmemMax;
    if (engine) {
        const size_t before = pd->launches.size();
        P.add_cols(direction == 1, outer, (int)n, inner, BUF_IN, BUF_OUT, scale);
        if (pd->launches.size() == before)   // n == 1 and nothing to do but copy / scale
            P.add_naive(MODE_NAIVE_COLS, outer * n * inner, 0, 1, inner, 1, direction == 1, BUF_IN, BUF_OUT, scale);
    } else {
        // The naive kernel re-reads whole columns while other threads store results, so it must never run with
        // src == dst.  `in` may equal `out` for this plan type: go through the workspace and copy back.
        pd->ws_bytes = (size_t)outer * n * inner * 8;
        P.add_naive(MODE_NAIVE_COLS, outer * n * inner, 0, (int)n, inner, (int)n, direction == 1, BUF_IN, BUF_WS, scale);
        P.add_naive(MODE_NAIVE_COLS, outer * n * inner, 0, 1, inner, 1, direction == 1, BUF_WS, BUF_OUT, 1.0f);
    }
    std::string d;
    for (size_t i = 0; i < pd->launches.size(); ++i) d += (i ? " ; " : "") + pd->launches[i].name;
    pd->description = d;
    return 0;
}

}  // namespace dft
