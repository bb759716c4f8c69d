// dft_fused2d.cuh -- shape-specialised fused rfft2 / irfft2 kernels (one image per tile).
//
// The whole image lives in shared memory exactly once: H x M complex for the forward
// transform (the real row of W = 2M floats *is* the packed complex row z[n] = x[2n] + i x[2n+1]),
// H x (M+1) complex for the inverse.  Row FFT length M = RA*RB, column FFT length H = CA*CB, each
// done as two in-place decimation-in-frequency steps with register butterflies.  All geometry is
// a template parameter, so every index is an immediate and every division a multiply-shift.
//
//   forward : TMA bulk load -> row step RA -> row step RB -> col step CA
//             -> [col step CB + real-FFT split + coalesced global stores]      (no separate store pass)
//   inverse : TMA bulk load -> col step CA -> [col step CB + Hermitian merge, rows to natural order]
//             -> row step RA -> [row step RB, natural order, dense pitch, 1/(H*W)] -> TMA bulk store
//
// The 2-D real-FFT split uses  X[q][k] = (Y[q][k] + conj Y[H-q][M-k])/2 - (i/2) w_W^k (Y[q][k] - conj Y[H-q][M-k])
// where Y is the full H x M complex 2-D FFT of the packed image; one thread owns both partner
// butterflies, so the pair never goes back through shared memory.
//
// Algorithmic HBM bytes per image: H*W*4 + H*(M+1)*8 (read once, written once); the kernels move
// exactly that.  3 CTAs are resident per SM (64.8 KB tile at 90x180), so one tile's load, another's
// butterflies and a third's stores overlap.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "dft_butterflies.cuh"
#include "dft_io16.cuh"

namespace dft {

template <int H_, int W_, int RA_, int RB_, int CA_, int CB_, int NT_, int MINB_>
struct FusedCfg {
    static constexpr int H = H_, W = W_, M = W_ / 2, RA = RA_, RB = RB_, CA = CA_, CB = CB_, NT = NT_, MINB = MINB_;
    static_assert(W_ % 2 == 0 && RA_ * RB_ == W_ / 2 && CA_ * CB_ == H_, "bad factorisation");
    static constexpr int NTW_R = (RA - 1) * RB;   // twR[(c-1)*RB + b] = w_M^(b*c)
    static constexpr int NTW_C = (CA - 1) * CB;   // twC[(c-1)*CB + b] = w_H^(b*c)
    static constexpr int KH = M / 2 + 1;          // column pairs (k, M-k), k = 0..M/2
    static constexpr int NTAB = NTW_R + NTW_C + KH;
    static constexpr int PI = M + 1;              // inverse pitch (one-sided spectrum row)
    static constexpr int SMEM_F = (H * M + NTAB) * 8 + 16;
    static constexpr int SMEM_I = (H * PI + NTAB) * 8 + 16;
    // Power-of-two chunk length: the forward kernel stores element e of chunk c at c*RB + (e ^ (c mod RB)).
    // Without it, lanes reading "element j of consecutive chunks" (row step RB, and the column-pair step
    // that walks consecutive frequencies k = c + RA*e) are RB * 8 bytes apart and collide 8-fold.
    static constexpr bool SWZ = (RB & (RB - 1)) == 0 && RB >= 8 && (NT_ % 32 == 0) && (32 % RB == 0);
    // Forward kernel, RB == 8 and rows a multiple of 16 elements long: rows r and r+1 start in the same bank, so
    // two rows' items in one half-warp collide 2-way.  Odd rows therefore swap the two 8-element halves of every
    // 16-element window (position ^ 8, i.e. chunk c sits at c ^ 1).  Column steps only ever combine rows of equal
    // parity at one position (CB even) or know the parity statically (the last step), so they follow for free.
    static constexpr bool RSW = SWZ && RB == 8 && (W_ / 2) % 16 == 0 && CB_ % 2 == 0 && H_ % 2 == 0;
    // Inverse kernel, RB == RA == 8: rows sit M+1 (odd) elements apart, so rows r and r+8 start 8 banks apart.
    // The row steps hand consecutive 8-lane groups the rows r, r+8, r+1, r+9, ... of each block of 16 rows: a
    // half-warp then covers all 16 banks (in natural row order two neighbouring rows overlap in 7 of 8).
    static constexpr bool IRW = SWZ && RB == 8 && RA == 8 && H_ % 16 == 0 && (NT_ / RA_) % 16 == 0;
    static constexpr bool BULK_REAL = (H * W * 4) % 16 == 0;
    static constexpr bool BULK_CPLX = (H * PI * 8) % 16 == 0;
};

#if defined(__CUDACC__)

__device__ __forceinline__ unsigned f2d_smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void f2d_mbar_init(unsigned bar)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// issue by one thread: 1-D TMA bulk copy global -> shared, completion on the mbarrier
__device__ __forceinline__ void f2d_bulk_load(void* dst, const void* src, unsigned bytes, unsigned bar)
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     f2d_smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}

__device__ __forceinline__ void f2d_mbar_wait(unsigned bar, unsigned phase)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "F2D_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra F2D_DONE;\n"
        "bra F2D_WAIT;\n"
        "F2D_DONE:\n"
        "}\n" ::"r"(bar),
        "r"(phase)
        : "memory");
}

// issue by one thread: 1-D TMA bulk copy shared -> global, then wait until the source was read
__device__ __forceinline__ void f2d_bulk_store_and_wait(void* gdst, const void* ssrc, unsigned bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(f2d_smem_u32(ssrc)),
                 "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

// ---- forward ---------------------------------------------------------------------------------
// kept-row index of spectrum row q for a truncated-mode plan (rows [0, kh) u [H - kh, H)), or -1
__device__ __forceinline__ int f2d_kept_row(int q, int H, int kh) { return q < kh ? q : (q >= H - kh ? q - (H - 2 * kh) : -1); }

// TR: truncated modes (trt_dft_plan_create_truncated) -- only rows [0, kh) u [H - kh, H), columns [0, kw) of the spectrum
// are stored, as a dense [2*kh][kw] block per image, and column pairs none of whose two columns is kept are not
// transformed in the last step at all.
//
// Both variants are ONE __global__ template: `Keep...` is empty for the full transform (the five-argument spec_fn signature)
// and <int, int> = (keep_h, keep_w) for truncated modes.  A shared __device__ body called from two thin kernels costs
// the inverse 60 bytes of spills under its register cap (ptxas 12.9; 0.041 -> 0.047 ms on 1024 x 128 x 128), a kernel
// that is its own entry point does not.
template <int I, class... Keep>
__device__ __forceinline__ int f2d_keep_arg(Keep... keep)
{
    const int v[] = {keep..., 0, 0};
    return v[I];
}

//
// IO = float, or __half / __nv_bfloat16 for typed-I/O plans: both tensors hold 16-bit elements, which are widened as the
// image is loaded into the (fp32) tile and rounded as the results are stored -- ordinary coalesced accesses in place of the
// bulk copies, everything in between unchanged.
template <class C, class IO, class... Keep>
__global__ void __launch_bounds__(C::NT, C::MINB)
    fused_rfft2_kernel(const float* __restrict__ in, float* __restrict__ out, long long nimg,
                       const float2* __restrict__ tables, float /*scale: forward is unnormalised*/, Keep... keep)
{
    constexpr bool TR = sizeof...(Keep) == 2;
    constexpr bool IO16 = !std::is_same<IO, float>::value;
    static_assert(TR || sizeof...(Keep) == 0, "Keep is () or (keep_h, keep_w)");
    static_assert(!(TR && IO16), "truncated-mode plans are fp32");
    const int kh = f2d_keep_arg<0>(keep...), kw = f2d_keep_arg<1>(keep...);
    constexpr int H = C::H, W = C::W, M = C::M, RA = C::RA, RB = C::RB, CA = C::CA, CB = C::CB, NT = C::NT;
    constexpr int KH = C::KH, MO = M + 1;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float2* buf = reinterpret_cast<float2*>(smem_raw);
    float2* twR = buf + H * M;
    float2* twC = twR + C::NTW_R;
    float2* twU = twC + C::NTW_C;            // 0.5 * w_W^k, k = 0..M/2
    const unsigned bar = f2d_smem_u32(twU + KH);
    const int tid = threadIdx.x;

    for (int i = tid; i < C::NTAB; i += NT) twR[i] = __ldg(tables + i);
    if (tid == 0) f2d_mbar_init(bar);
    __syncthreads();
    unsigned phase = 0;
    const bool bulk = !IO16 && C::BULK_REAL && (((uintptr_t)in & 15) == 0);
    const bool vec2 = (((uintptr_t)in & 7) == 0);

    for (long long img = blockIdx.x; img < nimg; img += gridDim.x) {
        const float* src = in + img * (long long)(H * W);
        if constexpr (IO16) {
            typedef typename Io16<IO>::v2 v2;
            const v2* s16 = reinterpret_cast<const v2*>(in) + img * (long long)(H * M);      // (4-byte aligned: W is even)
            for (int i = tid; i < H * M; i += NT) buf[i] = Io16<IO>::up(__ldg(s16 + i));
            __syncthreads();
        } else if (bulk) {
            if (tid == 0) f2d_bulk_load(buf, src, H * W * 4, bar);
            f2d_mbar_wait(bar, phase);
            phase ^= 1u;
        } else {
            if (vec2) {
                for (int i = tid; i < H * M; i += NT) buf[i] = __ldg(reinterpret_cast<const float2*>(src) + i);
            } else {
                float* bf = reinterpret_cast<float*>(buf);
                for (int i = tid; i < H * W; i += NT) bf[i] = __ldg(src + i);
            }
            __syncthreads();
        }

        // rows, step RA: elements b + a*RB of each row
#pragma unroll 1
        for (int t = tid; t < H * RB; t += NT) {
            const int row = t / RB, b = t - row * RB;
            float2* p = buf + row * M + b;
            float2 v[RA];
            if constexpr (C::RSW && RA % 2 == 0) {
                // the image arrives dense (rows r and r+1 share banks): odd rows read each pair (a, a+1) in the
                // opposite order, RB = 8 banks away from what the even row's lanes read at the same time
                const bool odd = (row & 1) != 0;
                const int o = odd ? RB : 0;
                static_for<0, RA / 2>([&](auto ac) {
                    constexpr int a = 2 * decltype(ac)::value;
                    const float2 x0 = p[a * RB + o], x1 = p[a * RB + RB - o];
                    v[a] = odd ? x1 : x0;
                    v[a + 1] = odd ? x0 : x1;
                });
            } else {
                static_for<0, RA>([&](auto ac) { constexpr int a = decltype(ac)::value; v[a] = p[a * RB]; });
            }
            dft_reg<RA>(v);
            static_for<1, RA>([&](auto cc) { constexpr int c = decltype(cc)::value; v[c] = cmul(v[c], twR[(c - 1) * RB + b]); });
            if constexpr (C::SWZ) {
                // swizzled store: not in place, but a row's RB items sit in one warp -> warp barrier between
                // the row's loads and its stores is enough
                __syncwarp();
                float2* prow = buf + row * M;
                const int rp = C::RSW ? (row & 1) : 0;
                static_for<0, RA>([&](auto cc) { constexpr int c = decltype(cc)::value; prow[(c ^ rp) * RB + (b ^ (c & (RB - 1)))] = v[c]; });
            } else {
                static_for<0, RA>([&](auto cc) { constexpr int c = decltype(cc)::value; p[c * RB] = v[c]; });
            }
        }
        __syncthreads();
        // rows, step RB: RB contiguous elements; row frequency k = k1 + RA*k2 ends at k1*RB + k2
#pragma unroll 1
        for (int t = tid; t < H * RA; t += NT) {
            const int k1 = t % RA;                                  // chunk index
            float2* p = buf + (C::RSW ? (t - k1 + (k1 ^ ((t / RA) & 1))) : t) * RB;
            const int sw = C::SWZ ? (k1 & (RB - 1)) : 0;
            float2 v[RB];
            static_for<0, RB>([&](auto jc) { constexpr int j = decltype(jc)::value; v[j] = p[j ^ sw]; });
            dft_reg<RB>(v);
            static_for<0, RB>([&](auto jc) { constexpr int j = decltype(jc)::value; p[j ^ sw] = v[j]; });
        }
        __syncthreads();
        // columns, step CA: rows b + a*CB of each column (threads run along the row: stride 1)
#pragma unroll 1
        for (int t = tid; t < CB * M; t += NT) {
            const int b = t / M;
            float2* p = buf + t;
            float2 v[CA];
            static_for<0, CA>([&](auto ac) { constexpr int a = decltype(ac)::value; v[a] = p[a * CB * M]; });
            dft_reg<CA>(v);
            static_for<1, CA>([&](auto cc) { constexpr int c = decltype(cc)::value; v[c] = cmul(v[c], twC[(c - 1) * CB + b]); });
            static_for<0, CA>([&](auto cc) { constexpr int c = decltype(cc)::value; p[c * CB * M] = v[c]; });
        }
        __syncthreads();
        // columns, step CB fused with the real-FFT split and the global stores
        float2* oimg = reinterpret_cast<float2*>(out) + img * (long long)(TR ? 2 * kh * kw : H * MO);
#pragma unroll 1
        for (int t = tid; t < CA * KH; t += NT) {
            const int q1 = t / KH, k = t - q1 * KH;
            if (TR && k >= kw && M - k >= kw) continue;          // neither column k nor column M - k is kept
            const int kk = (k == 0) ? 0 : M - k;
            const int swA = C::SWZ ? ((k % RA) & (RB - 1)) : 0, swB = C::SWZ ? ((kk % RA) & (RB - 1)) : 0;
            const int colA = (k % RA) * RB + ((k / RA) ^ swA);
            const int colB = (kk % RA) * RB + ((kk / RA) ^ swB);
            const int q1B = (q1 == 0) ? 0 : CA - q1;
            const float2* pa = buf + q1 * CB * M + colA;
            const float2* pb = buf + q1B * CB * M + colB;
            float2 A[CB], B[CB];
            // rows q1*CB + j and q1B*CB + j have parity j & 1 (CB even): odd rows hold the position ^ 8
            const int dA = C::RSW ? ((colA ^ 8) - colA) : 0, dB = C::RSW ? ((colB ^ 8) - colB) : 0;
            static_for<0, CB>([&](auto jc) {
                constexpr int j = decltype(jc)::value;
                A[j] = pa[j * M + (j & 1) * dA];
                B[j] = pb[j * M + (j & 1) * dB];
            });
            dft_reg<CB>(A);
            dft_reg<CB>(B);
            // A[k2] = Y[q1 + CA*k2][k]; its partner Y[(H-q) mod H][M-k] is B[CB-1-k2] (q1 > 0) or
            // B[(CB-k2) mod CB] (q1 == 0): rotate B by one in the latter case.
            if (q1 == 0) {
                const float2 b0 = B[0];
                static_for<0, CB - 1>([&](auto jc) { constexpr int j = decltype(jc)::value; B[j] = B[j + 1]; });
                B[CB - 1] = b0;
            }
            const float2 wh = twU[k];
            float2* oa = oimg + q1 * MO + k;                 // row q1 + CA*k2, column k
            float2* ob = oimg + (H - q1) * MO + (M - k);     // row H - q1 - CA*k2, column M-k
            static_for<0, CB>([&](auto kc) {
                constexpr int k2 = decltype(kc)::value;
                const float2 z1 = A[k2], z2 = B[CB - 1 - k2];
                const float ex = 0.5f * (z1.x + z2.x), ey = 0.5f * (z1.y - z2.y);
                const float dx = z1.x - z2.x, dy = z1.y + z2.y;
                const float tx = wh.x * dx - wh.y * dy, ty = wh.x * dy + wh.y * dx;
                if constexpr (IO16) {
                    typedef typename Io16<IO>::v2 v2;
                    v2* o16 = reinterpret_cast<v2*>(out) + img * (long long)(H * MO);
                    o16[(oa - oimg) + k2 * CA * MO] = Io16<IO>::down(float2{ex + ty, ey - tx});
                    long long ib = (ob - oimg) - k2 * CA * MO;
                    if (k2 == 0 && q1 == 0) ib -= H * MO;
                    o16[ib] = Io16<IO>::down(float2{ex - ty, -ey - tx});
                } else if constexpr (!TR) {
                    oa[k2 * CA * MO] = float2{ex + ty, ey - tx};
                    float2* pbk = ob - k2 * CA * MO;
                    if (k2 == 0 && q1 == 0) pbk -= H * MO;       // (H - 0) mod H = 0
                    *pbk = float2{ex - ty, -ey - tx};
                } else {
                    const int qa = q1 + CA * k2;
                    const int qb = (k2 == 0 && q1 == 0) ? 0 : H - q1 - CA * k2;
                    const int ia = f2d_kept_row(qa, H, kh), ib = f2d_kept_row(qb, H, kh);
                    if (k < kw && ia >= 0) oimg[ia * kw + k] = float2{ex + ty, ey - tx};
                    if (M - k < kw && ib >= 0) oimg[ib * kw + (M - k)] = float2{ex - ty, -ey - tx};
                }
            });
        }
        __syncthreads();
    }
}

// ---- inverse ---------------------------------------------------------------------------------
// TR: the input is the dense kept block [2*kh][kw]; it is expanded into the one-sided tile while it is loaded (zeros
// everywhere else), the rest is the ordinary kernel.
template <class C, class IO, class... Keep>
__global__ void __launch_bounds__(C::NT, C::MINB)
    fused_irfft2_kernel(const float* __restrict__ in, float* __restrict__ out, long long nimg,
                        const float2* __restrict__ tables, float scale /* 1/prod(signal dims) */, Keep... keep)
{
    constexpr bool TR = sizeof...(Keep) == 2;
    constexpr bool IO16 = !std::is_same<IO, float>::value;
    static_assert(TR || sizeof...(Keep) == 0, "Keep is () or (keep_h, keep_w)");
    static_assert(!(TR && IO16), "truncated-mode plans are fp32");
    const int kh = f2d_keep_arg<0>(keep...), kw = f2d_keep_arg<1>(keep...);
    constexpr int H = C::H, W = C::W, M = C::M, RA = C::RA, RB = C::RB, CA = C::CA, CB = C::CB, NT = C::NT;
    constexpr int KH = C::KH, PI = C::PI;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float2* buf = reinterpret_cast<float2*>(smem_raw);
    float2* twR = buf + H * PI;
    float2* twC = twR + C::NTW_R;
    float2* twU = twC + C::NTW_C;            // scale * w_W^k, k = 0..M/2
    const unsigned bar = f2d_smem_u32(twU + KH);
    const int tid = threadIdx.x;

    for (int i = tid; i < C::NTAB; i += NT) twR[i] = __ldg(tables + i);
    if (tid == 0) f2d_mbar_init(bar);
    __syncthreads();
    unsigned phase = 0;
    const bool bulk_in = !TR && !IO16 && C::BULK_CPLX && (((uintptr_t)in & 15) == 0);
    const bool bulk_out = !IO16 && C::BULK_REAL && (((uintptr_t)out & 15) == 0);

    for (long long img = blockIdx.x; img < nimg; img += gridDim.x) {
        const float2* src = reinterpret_cast<const float2*>(in) + img * (long long)(TR ? 2 * kh * kw : H * PI);
        if constexpr (TR) {
            // zero tile (128-bit stores), then the kept block scattered into it: cost follows the kept size
            static_assert((H * PI) % 2 == 0, "tile is cleared two elements at a time");
            const float4 z4 = float4{0.f, 0.f, 0.f, 0.f};
            for (int i = tid; i < H * PI / 2; i += NT) reinterpret_cast<float4*>(buf)[i] = z4;
            __syncthreads();
            const int total = 2 * kh * kw;
            for (int e = tid; e < total; e += NT) {
                const int qi = e / kw, k = e - qi * kw;
                buf[(qi < kh ? qi : qi + (H - 2 * kh)) * PI + k] = __ldg(src + e);
            }
            __syncthreads();
        } else if constexpr (IO16) {
            typedef typename Io16<IO>::v2 v2;
            const v2* s16 = reinterpret_cast<const v2*>(in) + img * (long long)(H * PI);
            for (int i = tid; i < H * PI; i += NT) buf[i] = Io16<IO>::up(__ldg(s16 + i));
            __syncthreads();
        } else if (bulk_in) {
            if (tid == 0) f2d_bulk_load(buf, src, H * PI * 8, bar);
            f2d_mbar_wait(bar, phase);
            phase ^= 1u;
        } else {
            for (int i = tid; i < H * PI; i += NT) buf[i] = __ldg(src + i);
            __syncthreads();
        }
        // columns, step CA over all M+1 columns
#pragma unroll 1
        for (int t = tid; t < CB * PI; t += NT) {
            const int b = t / PI;
            float2* p = buf + t;
            float2 v[CA];
            static_for<0, CA>([&](auto ac) { constexpr int a = decltype(ac)::value; v[a] = p[a * CB * PI]; });
            dft_reg<CA>(v);
            static_for<1, CA>([&](auto cc) { constexpr int c = decltype(cc)::value; v[c] = cmul(v[c], twC[(c - 1) * CB + b]); });
            static_for<0, CA>([&](auto cc) { constexpr int c = decltype(cc)::value; p[c * CB * PI] = v[c]; });
        }
        __syncthreads();
        // columns, step CB on the column pair (k, M-k), Hermitian merge, rows written in natural order.
        // Rounds hold whole column pairs (all CA row blocks), so read -> barrier -> write is race free.
        {
            constexpr int KPR = NT / CA;             // column pairs per round
            constexpr int NTR = KPR * CA;
#pragma unroll 1
            for (int k0 = 0; k0 < KH; k0 += KPR) {
                const int q1 = tid / KPR, kl = tid - q1 * KPR;       // lanes run along k: stride-1 shared-memory accesses
                const int k = k0 + kl;
                const bool active = (tid < NTR) && (k < KH);
                float2 A[CB], B[CB];
                if (active) {
                    const int kk = M - k;
                    const float2* pa = buf + q1 * CB * PI + k;
                    const float2* pb = buf + q1 * CB * PI + kk;
                    static_for<0, CB>([&](auto jc) { constexpr int j = decltype(jc)::value; A[j] = pa[j * PI]; B[j] = pb[j * PI]; });
                    dft_reg<CB>(A);
                    dft_reg<CB>(B);
                }
                __syncthreads();
                if (active) {
                    const int kk = M - k;
                    const float2 ws = twU[k];        // scale * (c, s) = scale * exp(-2 pi i k / W)
                    static_for<0, CB>([&](auto kc) {
                        constexpr int k2 = decltype(kc)::value;
                        // position q1*CB + k2 holds column frequency q1 + CA*k2 = inverse row (H - that) mod H
                        int r = H - q1 - CA * k2;
                        if (k2 == 0 && q1 == 0) r = 0;
                        const float2 a = A[k2], b = B[k2];
                        float2* row = buf + r * PI;
                        if (k == 0) {
                            row[0] = float2{scale * (a.x + b.x), scale * (a.x - b.x)};
                        } else {
                            const float sx = a.x + b.x, sy = a.y - b.y;
                            const float dx = a.x - b.x, dy = a.y + b.y;
                            const float ux = ws.x * dx + ws.y * dy, uy = ws.x * dy - ws.y * dx;
                            row[k] = float2{fmaf(scale, sx, -uy), fmaf(scale, sy, ux)};
                            if (kk != k) row[kk] = float2{fmaf(scale, sx, uy), fmaf(-scale, sy, ux)};
                        }
                    });
                }
            }
        }
        __syncthreads();
        // rows, step RA (pitch M+1)
#pragma unroll 1
        for (int t = tid; t < H * RB; t += NT) {
            const int g = t / RB, b = t - g * RB;
            const int row = C::IRW ? ((g & ~15) | ((g & 1) << 3) | ((g >> 1) & 7)) : g;
            float2* p = buf + row * PI + b;
            float2 v[RA];
            static_for<0, RA>([&](auto ac) { constexpr int a = decltype(ac)::value; v[a] = p[a * RB]; });
            dft_reg<RA>(v);
            static_for<1, RA>([&](auto cc) { constexpr int c = decltype(cc)::value; v[c] = cmul(v[c], twR[(c - 1) * RB + b]); });
            if constexpr (C::SWZ) {
                __syncwarp();                                  // same swizzle as the forward kernel
                float2* prow = buf + row * PI;
                static_for<0, RA>([&](auto cc) { constexpr int c = decltype(cc)::value; prow[c * RB + (b ^ (c & (RB - 1)))] = v[c]; });
            } else {
                static_for<0, RA>([&](auto cc) { constexpr int c = decltype(cc)::value; p[c * RB] = v[c]; });
            }
        }
        __syncthreads();
        // rows, step RB; results leave in natural order n = (M - f) mod M at dense pitch M.
        // Rounds hold whole rows in increasing order: new row r (pitch M) only overlaps old rows r-1, r.
        {
            constexpr int RPR = NT / RA;             // rows per round
            constexpr int NTR = RPR * RA;
#pragma unroll 1
            for (int r0 = 0; r0 < H; r0 += RPR) {
                const int g = tid / RA, k1 = tid - g * RA;
                const int rl = C::IRW ? ((g & ~15) | ((g & 1) << 3) | ((g >> 1) & 7)) : g;
                const int row = r0 + rl;
                const bool active = (tid < NTR) && (row < H);
                float2 v[RB];
                if (active) {
                    const float2* p = buf + row * PI + k1 * RB;
                    const int sw = C::SWZ ? (k1 & (RB - 1)) : 0;
                    static_for<0, RB>([&](auto jc) { constexpr int j = decltype(jc)::value; v[j] = p[j ^ sw]; });
                    dft_reg<RB>(v);
                }
                __syncthreads();
                if (active) {
                    float2* d = buf + row * M;
                    if constexpr (C::IRW) {
                        // the output rows are dense (pitch M: rows r and r+8 share banks), so the odd 8-lane groups
                        // store each pair (k2, k2+1) in the opposite order: their n differ by RA = 8 banks
                        const bool odd = (g & 1) != 0;
                        static_for<0, RB / 2>([&](auto kc) {
                            constexpr int k2 = 2 * decltype(kc)::value;
                            int n0 = M - k1 - RA * k2;
                            if (k2 == 0 && k1 == 0) n0 = 0;
                            const int n1 = M - k1 - RA * (k2 + 1);
                            const float2 a = odd ? v[k2 + 1] : v[k2], b = odd ? v[k2] : v[k2 + 1];
                            d[odd ? n1 : n0] = a;
                            d[odd ? n0 : n1] = b;
                        });
                    } else {
                        static_for<0, RB>([&](auto kc) {
                            constexpr int k2 = decltype(kc)::value;
                            int n = M - k1 - RA * k2;
                            if (k2 == 0 && k1 == 0) n = 0;
                            d[n] = v[k2];
                        });
                    }
                }
            }
        }
        float* dst = out + img * (long long)(H * W);
        if constexpr (IO16) {
            typedef typename Io16<IO>::v2 v2;
            v2* d16 = reinterpret_cast<v2*>(out) + img * (long long)(H * M);
            __syncthreads();
            for (int i = tid; i < H * M; i += NT) d16[i] = Io16<IO>::down(buf[i]);
            __syncthreads();
        } else if (bulk_out) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (tid == 0) f2d_bulk_store_and_wait(dst, buf, H * W * 4);
            if (!bulk_in) __syncthreads();
        } else {
            __syncthreads();
            for (int i = tid; i < H * M; i += NT) reinterpret_cast<float2*>(dst)[i] = buf[i];
            __syncthreads();
        }
    }
}

#endif  // __CUDACC__

}  // namespace dft
