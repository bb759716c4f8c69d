// dft_types.h -- plain structs shared by the host planner, the CUDA kernels and the
// test-only host emulator.  No CUDA runtime calls in here.
#pragma once
#include <stdint.h>
#include <vector_types.h>   // float2 (host-compatible header from the CUDA toolkit)

#if defined(__CUDACC__)
#define DFT_HD __host__ __device__ __forceinline__
#else
#define DFT_HD inline
#endif

namespace dft {

constexpr int kMaxSteps = 8;

// Radix schedule of one in-place, decimation-in-frequency complex FFT of length n.
// Step i works on blocks of length L_i = radix[i]*...*radix[nsteps-1]; frequency
// k = c_0 + r_0*(c_1 + r_1*(...)) ends at position sum_i c_i * (L_i / r_i).
struct LineFFT {
    int n;
    int nsteps;
    int radix[kMaxSteps];
    // ceil(2^32 / d) for the two per-step divisors (d = L_i/r_i and d = n/r_i); 0 encodes d == 1.
    // floor(x / d) == umulhi(x, magic) exactly whenever x * d < 2^32 (tile-local indices are < 2^16).
    unsigned mmag[kMaxSteps];
    unsigned bmag[kMaxSteps];
    // Row steps only (see row_step_order): bit 0 = consecutive work items walk the blocks (stride L) instead
    // of the offsets inside a block (mmag is then the magic of n/L); bit 1 = work items per line are padded to
    // a multiple of 16 so that no half-warp straddles two lines (bmag is then the magic of the padded count).
    unsigned char ord[kMaxSteps];
};

#if defined(__CUDACC__)
#define DFT_CXF __host__ __device__ constexpr
#else
#define DFT_CXF constexpr
#endif

// Shared-memory cost (wavefronts for 16 lines) of one radix-r row step on blocks of length L of n-point lines
// stored `pitch` complex elements apart, for work-item order `ord`: 16 consecutive work items form one
// 8-byte-access wavefront and every extra hit on one of the 16 bank pairs costs another.
DFT_CXF int row_step_cost(int n, int pitch, int L, int r, int ord)
{
    const int m = L / r, nblk = n / L, bpl = n / r;
    const int bplp = (ord & 2) ? (bpl + 15) / 16 * 16 : bpl;
    const int total = 16 * bplp;
    int cost = 0;
    for (int w0 = 0; w0 < total; w0 += 16) {
        int cnt[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
        int mx = 0;
        for (int l = 0; l < 16 && w0 + l < total; ++l) {
            const int t = w0 + l, line = t / bplp, u = t - line * bplp;
            if (u >= bpl) continue;
            int blk = u, b = 0;
            if (m > 1) {
                if (ord & 1) {
                    b = u / nblk;
                    blk = u - b * nblk;
                } else {
                    blk = u / m;
                    b = u - blk * m;
                }
            }
            const int bank = (line * pitch + blk * L + b) & 15;
            ++cnt[bank];
            if (cnt[bank] > mx) mx = cnt[bank];
        }
        cost += mx;
    }
    return cost;
}

// cheapest work-item order of a row step (0 keeps the plain order when nothing is gained)
DFT_CXF int row_step_order(int n, int pitch, int L, int r)
{
    int best = 0, bc = row_step_cost(n, pitch, L, r, 0);
    for (int o = 1; o < 4; ++o) {
        if ((o & 1) && (L / r <= 1 || n / L <= 1)) continue;
        const int c = row_step_cost(n, pitch, L, r, o);
        if (c < bc) {
            bc = c;
            best = o;
        }
    }
    return best;
}

DFT_CXF unsigned magic_u32(unsigned d) { return d <= 1u ? 0u : (unsigned)(((1ull << 32) + d - 1) / d); }

// apply row_step_order to every step of a row schedule (unskewed rows only)
DFT_CXF void set_row_orders(LineFFT& f, int pitch)
{
    int L = f.n;
    for (int i = 0; i < f.nsteps; ++i) {
        const int r = f.radix[i], m = L / r, bpl = f.n / r, nblk = f.n / L;
        const int o = row_step_order(f.n, pitch, L, r);
        f.ord[i] = (unsigned char)o;
        f.mmag[i] = magic_u32((unsigned)((o & 1) ? nblk : m));
        f.bmag[i] = magic_u32((unsigned)((o & 2) ? (bpl + 15) / 16 * 16 : bpl));
        L = m;
    }
}

enum Mode : int {
    MODE_F2D = 0,       // fused rfft2: real image [H][W] (W = 2M) -> complex [H][M+1]
    MODE_I2D = 1,       // fused irfft2
    MODE_F1D = 2,       // batched rows R2C, W = 2M
    MODE_I1D = 3,       // batched rows C2R, W = 2M
    MODE_F1D_ODD = 4,   // batched rows R2C, odd W (full complex FFT of zero-imag rows)
    MODE_I1D_ODD = 5,   // batched rows C2R, odd W (Hermitian extension, full complex FFT)
    MODE_C2C_COLS = 6,  // batched complex FFT along a strided axis: data[outer][N][inner]
    MODE_NAIVE_ROWS_R2C = 7,
    MODE_NAIVE_ROWS_C2R = 8,
    MODE_NAIVE_COLS = 9,
    MODE_SMALL_ROWS_R2C = 10,   // a few short rows (W = 2M <= 64): one thread per row, everything in registers
    MODE_SMALL_ROWS_C2R = 11,
};

// One kernel launch.  All pointers are device pointers (host pointers in the emulator).
struct KParams {
    int mode;
    int threads;             // CTA size
    int smem_bytes;          // dynamic shared memory
    long long ntiles;        // persistent loop bound
    long long nunits;        // images (2-D modes) / rows (1-D modes) / outer*chunks (cols)
    int G;                   // units per tile (2-D: images, 1-D: rows); cols: columns per tile (TM)
    int H;                   // 2-D: rows per image; cols mode: FFT length N
    int M;                   // complex row-FFT length (W/2 for even W, W for odd W)
    int W;                   // real row length
    int P;                   // shared-memory pitch in complex elements
    LineFFT row;             // FFT along the contiguous axis (length M)
    LineFFT col;             // FFT along the strided axis (length H)
    const float2* tw_row;    // w_M^j,  j in [0,M)        w_n = exp(-2*pi*i/n)
    const float2* tw_row_t;  // row twiddles per step, transposed: step s (block length L, radix r, m = L/r) holds
                             // w_L^(b*c) at [(c-1)*m + b]; consecutive work items b read consecutive entries.
                             // Steps follow each other (n - 1 entries in all).  Null: use tw_row.
    const float2* tw_col;    // w_H^j,  j in [0,H)
    const float2* tw_unt;    // w_W^k,  k in [0,M]        (real-FFT split twiddles)
    const int* perm_row;     // smem position of output index k of the row FFT
    const int* perm_col;     // smem position of output index q of the column FFT
    float scale;             // applied when the result is stored
    long long inner;         // cols mode: contiguous inner extent (complex elements)
    long long outer;         // cols mode: number of [N][inner] slabs
    long long chunks;        // cols mode: column chunks per slab = ceil(inner / G)
    int pad;                 // rows modes: 1 = power-of-two rows stored skewed in shared memory,
                             // physical index p + p/16 + p/256 (kills the 16-way bank conflicts)
    // MODE_F1D only: write the result column-split for an all-to-all (pencil decomposition).  Rows come in
    // images of pack_hn rows; within an image, columns [pack_c0[j], pack_c0[j+1]) of all its rows form the
    // contiguous block destined for rank j.  pack_n == 0: ordinary [rows][M+1] output.
    int pack_n;
    int pack_c0[9];
    int pack_hn;
    // Peer variant of the same (multi-GPU pencil rfft2, exchange fused into the row pass): block j does not go into
    // `out` but straight into rank j's column slab [image][pack_H rows][cn_j columns] at pack_peer[j] (a pointer into
    // that GPU's memory, written with ordinary stores); this rank's rows start at row pack_h0 of every image.
    // pack_peer_on == 0: blocks go into `out` as described above.
    int pack_peer_on;
    int pack_H;
    int pack_h0;
    float2* pack_peer[8];
    int pack_pitch[8];       // row pitch of rank j's slab in complex elements (>= its column count)
    // pack_bulk != 0 (pipelined kernel only): the split result of a tile is staged in shared memory in natural order
    // (pack_stage_off bytes into the CTA's shared memory, pack_stage_bytes per compute group, row pitch pack_mo_pad
    // elements) and every (row, destination) run leaves as ONE bulk copy (cp.async.bulk shared -> peer global), which
    // keeps the NVLink stores going while the group transforms its next tile.  Needs even column starts and even pitches.
    int pack_bulk;
    int pack_stage_off;
    int pack_stage_bytes;
    int pack_mo_pad;
    // MODE_C2C_COLS, mirror image of the above (pencil irfft2, exchange fused into the column pass): output row q of
    // every image belongs to the rank j with split_r0[j] <= q < split_r0[j+1] and is stored into that rank's row band
    // [image][rows of rank j][split_pitch columns] at split_peer[j], starting at column split_c0 (this rank's first
    // column).  split_n == 0: rows go into `out`.
    int split_n;
    int split_r0[9];
    int split_pitch;
    int split_c0;
    // When this launch is the second pass of a four-step FFT it sees the tensor as [outer][N1][N2 * inner]: its row q'
    // and column j' are row q' * split_n2 + j' / split_inner, column j' % split_inner of the image.  Ordinary pass:
    // split_n2 == 1.
    int split_n2;
    int split_inner;
    float2* split_peer[8];
    // MODE_C2C_COLS as the first pass of a four-step FFT of a long axis Ntot = n1split * N:
    // slab o_ = (o_hi, n1) gathers rows n1 + n1split*n2 (n2 = 0..N-1) of image o_hi, stores them as the
    // contiguous rows n1*N + k2 and multiplies row k2 by w_Ntot^(n1*k2) (tw_four, Ntot entries).
    // n1split <= 1: ordinary pass.  rev_in: read the axis index-reversed, row (Ntot - n) mod Ntot for n,
    // which turns the forward transform into the inverse one (IDFT(x)[n] = DFT(x reversed)[n]).
    int n1split;
    const float2* tw_four;
    int rev_in;
    // Four-step passes through a PADDED workspace (row pitch a multiple of 16 bytes, so that pass 2 can fetch its tiles
    // with 2-D TMA tensor copies):
    //   out_rowlen : pass 1 -- pitch of the rows it writes, in complex elements (0: = inner)
    //   unflat_in  : pass 2 -- it sees the workspace as [outer][N1][N2 * unflat_in]; column j' of its tile is column
    //                j' % unflat_in (dropped when >= unflat_w) of workspace row k2 = j' / unflat_in, and the result goes to
    //                row q' * N2 + k2 of the image, pitch unflat_w.  0: input and output are flattened alike.
    //   tma_cols   : pass 2 -- the executor provides a tensor map over the workspace; tiles arrive by one
    //                cp.async.bulk.tensor.2d each instead of per-element copies
    int out_rowlen;
    int unflat_in;
    int unflat_w;
    int tma_cols;
    int inverse;             // cols mode: 1 = inverse transform (output index reversed)
    int scratch_off;         // != 0: the tile is followed (scratch_off complex elements on) by a second copy of its
                             // footprint, the out-of-place target of the generic prime-radix steps (radices above 31)
    int in_place_ok;         // informational
    int io_align;            // 16 if in/out base pointers and unit strides allow 16-byte bulk copies
};

}  // namespace dft
