// dft_pipe2d.cuh -- persistent, warp-specialised fused rfft2 / irfft2 kernels.
//
// One CTA per SM.  A producer warp streams whole images into a ring of NS shared-memory stages
// with 1-D TMA bulk copies (cp.async.bulk + mbarrier full/empty pairs); NG compute groups of GT
// threads each take every NG-th image, transform it in place in its stage and hand the stage
// back.  Because one stage is always being filled while the groups work on the others, HBM
// reads run continuously under the butterflies; groups synchronise among themselves with named
// barriers (bar.sync id, GT), never CTA-wide.
//
//   forward : rows DIF  [radix RA, stride RB, twiddle] -> [radix RB, contiguous 128-bit chunks]
//             cols DIF  [radix CA, stride CB rows, twiddle] -> [radix CB on the column pair
//             (k, M-k) + real-FFT split + coalesced 8-byte global stores]
//   inverse : cols DIF  [radix CA] -> [radix CB on the pair (k, M-k) + Hermitian merge; rows land
//             in natural order, columns in digit-reversed order, pitch M, conjugated, scaled]
//             rows DIT  [radix RB chunks, twiddle] -> [radix RA, stride RB -> natural order,
//             conjugated back] -> one TMA bulk store of the dense real image
//
// Shared-memory bank behaviour (8-byte elements, 16 lanes per wavefront): a strided step maps
// consecutive work items t to addresses congruent to t mod 16 when RB == pitch mod 16
// (90 = 5*16 + 10, RB = 10), and the contiguous chunks are read with 128-bit accesses whose
// 80-byte lane stride is odd in 16-byte units -- both conflict free.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "dft_butterflies.cuh"

namespace dft {

template <int H_, int W_, int RA_, int RB_, int CA_, int CB_, int GT_, int NG_, int NS_>
struct PipeCfg {
    static constexpr int H = H_, W = W_, M = W_ / 2, RA = RA_, RB = RB_, CA = CA_, CB = CB_;
    static constexpr int GT = GT_, NG = NG_, NS = NS_;
    static_assert(W_ % 2 == 0 && RA_ * RB_ == W_ / 2 && CA_ * CB_ == H_, "bad factorisation");
    static_assert(RB_ % 2 == 0, "row chunks are accessed 128 bits at a time");
    static constexpr int PI = M + 1;
    static constexpr int KH = M / 2 + 1;
    // the inverse kernel's merge step needs every (k, M-k) item in one round of a group
    static constexpr bool kInverseOk = CA_ * KH <= GT_;
    static constexpr int TILE_F = H * M * 8;       // bytes: real image == packed complex image
    static constexpr int TILE_I = H * PI * 8;      // bytes: one-sided spectrum
    static_assert(TILE_F % 16 == 0 && TILE_I % 16 == 0, "bulk copies need 16-byte multiples");
    static constexpr int STAGE = (TILE_I + 127) / 128 * 128;
    static constexpr int NTW_RF = (RA - 1) * RB;   // forward rows: twR[(c-1)*RB + b]   = w_M^(b c)
    static constexpr int NTW_RI = (RB - 1) * RA;   // inverse rows: twR[(n2-1)*RA + k1] = w_M^(k1 n2)
    static constexpr int NTW_C = (CA - 1) * CB;    // columns:      twC[(c-1)*CB + b]   = w_H^(b c)
    static constexpr int NTAB_F = NTW_RF + NTW_C + KH;
    static constexpr int NTAB_I = NTW_RI + NTW_C + KH;
    static constexpr int NTAB = NTAB_F > NTAB_I ? NTAB_F : NTAB_I;
    static constexpr int THREADS = NG * GT + 32;
    static constexpr int SMEM = NS * STAGE + NTAB * 8 + 2 * NS * 8 + 16;
};

#if defined(__CUDACC__)

namespace pipe {

__device__ __forceinline__ unsigned s32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "PIPE_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra PIPE_DONE;\n"
        "bra PIPE_WAIT;\n"
        "PIPE_DONE:\n"
        "}\n" ::"r"(bar),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ bool mbar_test(unsigned bar, unsigned parity)   // non-blocking
{
    unsigned ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_arrive(unsigned bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void group_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// producer lane: keep the ring of stages full
template <class C>
__device__ __forceinline__ void producer(const unsigned char* __restrict__ src, long long tile_bytes_stride, unsigned bytes,
                                         long long nimg, unsigned char* stages, unsigned full0, unsigned empty0)
{
    long long it = 0;
    for (long long tile = blockIdx.x; tile < nimg; tile += gridDim.x, ++it) {
        const int s = (int)(it % C::NS);
        if (it >= C::NS) mbar_wait(empty0 + 8 * s, (unsigned)(((it / C::NS) - 1) & 1));
        fence_async();
        const unsigned bar = full0 + 8 * s;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         s32(stages + (size_t)s * C::STAGE)),
                     "l"(src + tile * tile_bytes_stride), "r"(bytes), "r"(bar)
                     : "memory");
    }
}

template <int R>
__device__ __forceinline__ void load_chunk(float2 (&v)[R], const float2* p)
{
    static_assert(R % 2 == 0, "");
    const float4* p4 = reinterpret_cast<const float4*>(p);
    static_for<0, R / 2>([&](auto jc) {
        constexpr int j = decltype(jc)::value;
        const float4 q = p4[j];
        v[2 * j] = float2{q.x, q.y};
        v[2 * j + 1] = float2{q.z, q.w};
    });
}
template <int R>
__device__ __forceinline__ void store_chunk(const float2 (&v)[R], float2* p)
{
    float4* p4 = reinterpret_cast<float4*>(p);
    static_for<0, R / 2>([&](auto jc) {
        constexpr int j = decltype(jc)::value;
        p4[j] = float4{v[2 * j].x, v[2 * j].y, v[2 * j + 1].x, v[2 * j + 1].y};
    });
}

}  // namespace pipe

// One item of the forward kernel's last phase: radix-CB butterflies of column k (row block q1) and of
// its partner column M-k (row block CA-q1), the real-FFT split of the CB partner pairs, and the
// 2*CB coalesced 8-byte stores.  A[k2] = Y[q1 + CA*k2][k]; its partner Y[(H-q) mod H][M-k] is
// B[CB-1-k2] for q1 > 0 and B[(CB-k2) mod CB] for q1 == 0 (Q0).
template <class C, bool Q0>
__device__ __forceinline__ void split_store(const float2* buf, const float2* twU, float2* oimg, int q1, int k)
{
    constexpr int H = C::H, M = C::M, RA = C::RA, RB = C::RB, CA = C::CA, CB = C::CB, MO = M + 1;
    const int kk = (k == 0) ? 0 : M - k;
    const int colA = (k % RA) * RB + k / RA;
    const int colB = (kk % RA) * RB + kk / RA;
    const int q1B = Q0 ? 0 : CA - q1;
    const float2* pa = buf + q1 * CB * M + colA;
    const float2* pb = buf + q1B * CB * M + colB;
    float2 A[CB], B[CB];
    static_for<0, CB>([&](auto jc) { constexpr int j = decltype(jc)::value; A[j] = pa[j * M]; B[j] = pb[j * M]; });
    dft_reg<CB>(A);
    dft_reg<CB>(B);
    const float2 wh = twU[k];                            // 0.5 * w_W^k
    float2* oa = oimg + q1 * MO + k;                     // row q1 + CA*k2, column k
    float2* ob = oimg + (H - q1) * MO + (M - k);         // row H - q1 - CA*k2, column M-k
    static_for<0, CB>([&](auto kc) {
        constexpr int k2 = decltype(kc)::value;
        constexpr int kp = Q0 ? (CB - k2) % CB : CB - 1 - k2;
        const float2 z1 = A[k2], z2 = B[kp];
        const float sx = z1.x + z2.x, sy = z1.y - z2.y;  // z1 + conj z2
        const float dx = z1.x - z2.x, dy = z1.y + z2.y;  // z1 - conj z2
        const float tx = wh.x * dx - wh.y * dy, ty = wh.x * dy + wh.y * dx;
        oa[k2 * CA * MO] = float2{fmaf(0.5f, sx, ty), fmaf(0.5f, sy, -tx)};
        float2* pbk = ob - k2 * CA * MO;
        if (Q0 && k2 == 0) pbk -= H * MO;                // (H - 0) mod H = 0
        *pbk = float2{fmaf(0.5f, sx, -ty), fmaf(-0.5f, sy, -tx)};
    });
}

// ---- forward ---------------------------------------------------------------------------------
template <class C>
__global__ void __launch_bounds__(C::THREADS, 1)
    pipe_rfft2_kernel(const float* __restrict__ in, float* __restrict__ out, long long nimg,
                      const float2* __restrict__ tables, float /*scale*/)
{
    constexpr int H = C::H, M = C::M, RA = C::RA, RB = C::RB, CA = C::CA, CB = C::CB, GT = C::GT, NG = C::NG, NS = C::NS;
    constexpr int KH = C::KH, MO = M + 1;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned char* stages = smem_raw;
    float2* twR = reinterpret_cast<float2*>(smem_raw + NS * C::STAGE);
    float2* twC = twR + C::NTW_RF;
    float2* twU = twC + C::NTW_C;                       // 0.5 * w_W^k
    const unsigned full0 = pipe::s32(twR + C::NTAB);
    const unsigned empty0 = full0 + 8 * NS;
    const int tid = threadIdx.x;

    for (int i = tid; i < C::NTAB_F; i += C::THREADS) twR[i] = __ldg(tables + i);
    if (tid == 0) {
        for (int s = 0; s < NS; ++s) {
            pipe::mbar_init(full0 + 8 * s, 1);
            pipe::mbar_init(empty0 + 8 * s, GT / 32);   // every warp of the owning group releases the stage
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        pipe::fence_async();
    }
    __syncthreads();

    if (tid >= NG * GT) {                               // producer warp
        if (tid == NG * GT)
            pipe::producer<C>(reinterpret_cast<const unsigned char*>(in), (long long)C::TILE_F, C::TILE_F, nimg, stages, full0, empty0);
        return;
    }
    const int g = tid / GT;
    const int gt = tid - g * GT;
    const int bar_id = 1 + g;

    long long it = g;
    for (long long tile = blockIdx.x + (long long)g * gridDim.x; tile < nimg; tile += (long long)NG * gridDim.x, it += NG) {
        const int s = (int)(it % NS);
        float2* buf = reinterpret_cast<float2*>(stages + (size_t)s * C::STAGE);
        pipe::mbar_wait(full0 + 8 * s, (unsigned)((it / NS) & 1));

        // rows, DIF step RA: elements b + a*RB
#pragma unroll 2
        for (int t = gt; t < H * RB; t += GT) {
            const int row = t / RB, b = t - row * RB;
            float2* p = buf + row * M + b;
            float2 v[RA];
            static_for<0, RA>([&](auto ac) { constexpr int a = decltype(ac)::value; v[a] = p[a * RB]; });
            dft_reg<RA>(v);
            static_for<1, RA>([&](auto cc) { constexpr int c = decltype(cc)::value; v[c] = cmul(v[c], twR[(c - 1) * RB + b]); });
            static_for<0, RA>([&](auto cc) { constexpr int c = decltype(cc)::value; p[c * RB] = v[c]; });
        }
        pipe::group_sync(bar_id, GT);
        // rows, DIF step RB: contiguous chunk; frequency k = k1 + RA*k2 stays at k1*RB + k2
#pragma unroll 2
        for (int t = gt; t < H * RA; t += GT) {
            float2* p = buf + t * RB;
            float2 v[RB];
            pipe::load_chunk<RB>(v, p);
            dft_reg<RB>(v);
            pipe::store_chunk<RB>(v, p);
        }
        pipe::group_sync(bar_id, GT);
        // columns, DIF step CA: rows b + a*CB; threads run along the row
#pragma unroll 2
        for (int t = gt; t < CB * M; t += GT) {
            const int b = t / M;
            float2* p = buf + t;
            float2 v[CA];
            static_for<0, CA>([&](auto ac) { constexpr int a = decltype(ac)::value; v[a] = p[a * CB * M]; });
            dft_reg<CA>(v);
            static_for<1, CA>([&](auto cc) { constexpr int c = decltype(cc)::value; v[c] = cmul(v[c], twC[(c - 1) * CB + b]); });
            static_for<0, CA>([&](auto cc) { constexpr int c = decltype(cc)::value; p[c * CB * M] = v[c]; });
        }
        pipe::group_sync(bar_id, GT);
        // columns, step CB on the pair (k, M-k) + real-FFT split + global stores.  Items with q1 == 0
        // pair differently (see split_store), so they get whole warps of their own: threads [0, KHP)
        // take q1 == 0 (KHP = KH rounded up to a warp), the rest take q1 = 1 .. CA-1.
        {
            constexpr int KHP = (KH + 31) / 32 * 32;
            float2* oimg = reinterpret_cast<float2*>(out) + tile * (long long)(H * MO);
#pragma unroll 1
            for (int t = gt; t < KHP + (CA - 1) * KH; t += GT) {
                if (t < KHP) {
                    if (t < KH) split_store<C, true>(buf, twU, oimg, 0, t);
                } else {
                    const int u = t - KHP;
                    const int q1 = 1 + u / KH, k = u - (q1 - 1) * KH;
                    split_store<C, false>(buf, twU, oimg, q1, k);
                }
            }
        }
        // release: every warp hands the stage back on its own (no group barrier at the tile end)
        pipe::fence_async();
        __syncwarp();
        if ((gt & 31) == 0) pipe::mbar_arrive(empty0 + 8 * s);
    }
}

// ---- inverse ---------------------------------------------------------------------------------
template <class C>
__global__ void __launch_bounds__(C::THREADS, 1)
    pipe_irfft2_kernel(const float* __restrict__ in, float* __restrict__ out, long long nimg,
                       const float2* __restrict__ tables, float scale)
{
    static_assert(C::kInverseOk, "inverse kernel: group too small for the one-round merge step");
    constexpr int H = C::H, M = C::M, RA = C::RA, RB = C::RB, CA = C::CA, CB = C::CB, GT = C::GT, NG = C::NG, NS = C::NS;
    constexpr int KH = C::KH, PI = C::PI;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned char* stages = smem_raw;
    float2* twR = reinterpret_cast<float2*>(smem_raw + NS * C::STAGE);
    float2* twC = twR + C::NTW_RI;
    float2* twU = twC + C::NTW_C;                       // scale * w_W^k
    const unsigned full0 = pipe::s32(reinterpret_cast<float2*>(smem_raw + NS * C::STAGE) + C::NTAB);
    const unsigned empty0 = full0 + 8 * NS;
    const int tid = threadIdx.x;

    for (int i = tid; i < C::NTAB_I; i += C::THREADS) twR[i] = __ldg(tables + i);
    if (tid == 0) {
        for (int s = 0; s < NS; ++s) {
            pipe::mbar_init(full0 + 8 * s, 1);
            pipe::mbar_init(empty0 + 8 * s, 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        pipe::fence_async();
    }
    __syncthreads();

    if (tid >= NG * GT) {
        if (tid == NG * GT)
            pipe::producer<C>(reinterpret_cast<const unsigned char*>(in), (long long)C::TILE_I, C::TILE_I, nimg, stages, full0, empty0);
        return;
    }
    const int g = tid / GT;
    const int gt = tid - g * GT;
    const int bar_id = 1 + g;

    long long it = g;
    for (long long tile = blockIdx.x + (long long)g * gridDim.x; tile < nimg; tile += (long long)NG * gridDim.x, it += NG) {
        const int s = (int)(it % NS);
        float2* buf = reinterpret_cast<float2*>(stages + (size_t)s * C::STAGE);
        pipe::mbar_wait(full0 + 8 * s, (unsigned)((it / NS) & 1));

        // columns, DIF step CA over all M+1 columns (pitch M+1)
#pragma unroll 1
        for (int t = gt; t < CB * PI; t += GT) {
            const int b = t / PI;
            float2* p = buf + t;
            float2 v[CA];
            static_for<0, CA>([&](auto ac) { constexpr int a = decltype(ac)::value; v[a] = p[a * CB * PI]; });
            dft_reg<CA>(v);
            static_for<1, CA>([&](auto cc) { constexpr int c = decltype(cc)::value; v[c] = cmul(v[c], twC[(c - 1) * CB + b]); });
            static_for<0, CA>([&](auto cc) { constexpr int c = decltype(cc)::value; p[c * CB * PI] = v[c]; });
        }
        pipe::group_sync(bar_id, GT);
        // columns, step CB on the pair (k, M-k) + Hermitian merge.  One round holds every item, so
        // read -> barrier -> write may re-pitch the tile: row r (natural order) at r*M, column
        // frequency k at (k mod RA)*RB + k/RA, value conj(Z'[k]) * scale.
        {
            // items: threads [0, CA*(KH-1)) take (q1, k >= 1); the CA items with k == 0 (which pair column 0
            // with the Nyquist column M and ignore imaginary parts) sit together at the end, so that at most
            // one warp sees both kinds.
            constexpr int NK = KH - 1;
            const bool active = gt < CA * KH;
            const bool k_is_0 = gt >= CA * NK;
            const int q1 = k_is_0 ? gt - CA * NK : gt / NK;
            const int k = k_is_0 ? 0 : 1 + (gt - q1 * NK);
            const int kk = M - k;
            float2 A[CB], B[CB];
            if (active) {
                const float2* pa = buf + q1 * CB * PI + k;
                const float2* pb = buf + q1 * CB * PI + kk;
                static_for<0, CB>([&](auto jc) { constexpr int j = decltype(jc)::value; A[j] = pa[j * PI]; B[j] = pb[j * PI]; });
                dft_reg<CB>(A);
                dft_reg<CB>(B);
            }
            pipe::group_sync(bar_id, GT);
            if (active) {
                // column frequency q1 + CA*k2 is inverse row (H - q1 - CA*k2) mod H: a fixed pointer minus
                // compile-time offsets (the one wrap, q1 == 0 && k2 == 0 -> row 0, is patched below)
                if (k_is_0) {
                    float2* p0 = buf + (H - q1) * M;
                    static_for<0, CB>([&](auto kc) {
                        constexpr int k2 = decltype(kc)::value;
                        float2* w0 = p0 - k2 * CA * M;
                        if (k2 == 0 && q1 == 0) w0 -= H * M;
                        *w0 = float2{scale * (A[k2].x + B[k2].x), scale * (B[k2].x - A[k2].x)};
                    });
                } else {
                    const float2 ws = twU[k];
                    const int posA = (k % RA) * RB + k / RA;
                    const int posB = (kk % RA) * RB + kk / RA;
                    float2* pA = buf + (H - q1) * M + posA;
                    float2* pB = buf + (H - q1) * M + posB;
                    static_for<0, CB>([&](auto kc) {
                        constexpr int k2 = decltype(kc)::value;
                        const float2 a = A[k2], b = B[k2];
                        float2* wa = pA - k2 * CA * M;
                        float2* wb = pB - k2 * CA * M;
                        if (k2 == 0 && q1 == 0) {
                            wa -= H * M;
                            wb -= H * M;
                        }
                        const float sx = a.x + b.x, sy = a.y - b.y;
                        const float dx = a.x - b.x, dy = a.y + b.y;
                        const float ux = ws.x * dx + ws.y * dy, uy = ws.x * dy - ws.y * dx;
                        // Z'[k] = (sx - uy, sy + ux), Z'[M-k] = (sx + uy, -sy + ux); stored conjugated.
                        // (k == M-k, M even: both stores write the same value to the same place.)
                        *wa = float2{fmaf(scale, sx, -uy), -fmaf(scale, sy, ux)};
                        *wb = float2{fmaf(scale, sx, uy), fmaf(scale, sy, -ux)};
                    });
                }
            }
        }
        pipe::group_sync(bar_id, GT);
        // rows, DIT step RB: contiguous chunk k1*RB + k2 -> n2, twiddle w_M^(k1 n2)
#pragma unroll 1
        for (int t = gt; t < H * RA; t += GT) {
            const int row = t / RA, k1 = t - row * RA;
            float2* p = buf + t * RB;
            float2 v[RB];
            pipe::load_chunk<RB>(v, p);
            dft_reg<RB>(v);
            static_for<1, RB>([&](auto nc) { constexpr int n2 = decltype(nc)::value; v[n2] = cmul(v[n2], twR[(n2 - 1) * RA + k1]); });
            pipe::store_chunk<RB>(v, p);
        }
        pipe::group_sync(bar_id, GT);
        // rows, DIT step RA: elements k1*RB + n2 -> n = n1*RB + n2 (natural order); conjugate back
#pragma unroll 1
        for (int t = gt; t < H * RB; t += GT) {
            const int row = t / RB, n2 = t - row * RB;
            float2* p = buf + row * M + n2;
            float2 v[RA];
            static_for<0, RA>([&](auto ac) { constexpr int a = decltype(ac)::value; v[a] = p[a * RB]; });
            dft_reg<RA>(v);
            static_for<0, RA>([&](auto cc) { constexpr int c = decltype(cc)::value; p[c * RB] = float2{v[c].x, -v[c].y}; });
        }
        pipe::fence_async();
        pipe::group_sync(bar_id, GT);
        if (gt == 0) {
            float* dst = out + tile * (long long)(H * C::W);
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(pipe::s32(buf)),
                         "r"((unsigned)C::TILE_F)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            pipe::mbar_arrive(empty0 + 8 * s);
        }
    }
    // the last bulk stores must have completed before the CTA (and its shared memory) goes away
    if (gt == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

#endif  // __CUDACC__

}  // namespace dft
