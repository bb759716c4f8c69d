// dft_kernels.cu -- sm_100a kernels, the plan executor and the C ABI of include/trt_dft.h.
//
// One persistent kernel template per engine mode: each CTA loops over tiles; a tile is brought
// into shared memory (cp.async.bulk + mbarrier when the source is 16-byte aligned, vectorised
// loads otherwise), transformed in place by the phases of dft_programs.cuh, and written back
// with coalesced stores.  Several CTAs are resident per SM, so one tile's load overlaps
// another tile's butterflies and a third one's stores.
//
// There is no CPU route in this file: if a kernel cannot be launched the call returns an error.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <new>
#include <vector>

#include "../../include/trt_dft.h"
#include "dft_spec.h"
#include "dft_launch.h"
#include "dft_planner.h"
#include "dft_programs.cuh"

namespace {

using dft::KParams;
using dft::kernel_fn;
using dft::pipe_fn;
using dft::kPipeThreads;

using dft::SpecEntry;
using dft::spec_fn;
static const SpecEntry* find_spec(int H, int W) { return dft::find_spec(H, W); }

// Grid of a shape-specialised kernel over `nimg` images with `occ` CTAs resident per SM.  The register-resident kernels are
// persistent by construction (one CTA per SM, images handed round in a fixed order).  The 3-CTA-per-SM fused kernels loop
// over images too, but once the batch is more than two full rounds they run one CTA per image instead: the hardware then
// hands images to whichever SM frees up first, which measured 5-10 % faster than the fixed round-robin (128 x 128: 1024
// images 0.0369 -> 0.0350 ms, 2048 images 0.068 -> 0.062, 16384 images 0.453 -> 0.430), while just above one round the
// CTAs' set-up costs more than the balance returns (500 images 0.0190 vs 0.0205 ms).  TRT_DFT_SPEC_PERSISTENT=1: A/B.
static int spec_grid_for(const SpecEntry& e, long long nimg, int num_sms, int occ)
{
    const long long slots = (long long)num_sms * occ;
    if (!e.pipelined && nimg > 2 * slots && nimg <= 0x7fffffffLL && !getenv("TRT_DFT_SPEC_PERSISTENT")) return (int)nimg;
    return (int)std::min<long long>(nimg, slots);
}


// twR[(c-1)*RB + b] = w_M^(b c) | twC[(c-1)*CB + b] = w_H^(b c) | twU[k] = f * w_W^k, k = 0..M/2
std::vector<float2> spec_tables(const SpecEntry& e, bool inverse, double scale)
{
    const int M = e.W / 2;
    const double two_pi = 6.283185307179586476925286766559;
    std::vector<float2> t;
    if (e.pipelined && inverse) {
        // inverse rows run decimation-in-time: twR[(n2-1)*RA + k1] = w_M^(k1 n2)
        for (int n2 = 1; n2 < e.RB; ++n2)
            for (int k1 = 0; k1 < e.RA; ++k1) {
                const double a = two_pi * (double)((k1 * n2) % M) / M;
                t.push_back(float2{(float)cos(a), (float)-sin(a)});
            }
    } else {
        for (int c = 1; c < e.RA; ++c)
            for (int b = 0; b < e.RB; ++b) {
                const double a = two_pi * (double)((b * c) % M) / M;
                t.push_back(float2{(float)cos(a), (float)-sin(a)});
            }
    }
    for (int c = 1; c < e.CA; ++c)
        for (int b = 0; b < e.CB; ++b) {
            const double a = two_pi * (double)((b * c) % e.H) / e.H;
            t.push_back(float2{(float)cos(a), (float)-sin(a)});
        }
    const double f = inverse ? scale : 0.5;   // inverse: the plan's 1/prod(dims) rides on the merge twiddle
    for (int k = 0; k <= M / 2; ++k) {
        const double a = two_pi * (double)k / e.W;
        t.push_back(float2{(float)(f * cos(a)), (float)(-f * sin(a))});
    }
    return t;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda)
typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
encode_tiled_fn tensor_map_encoder()
{
    static encode_tiled_fn fn = []() -> encode_tiled_fn {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            return nullptr;
        }
        return (encode_tiled_fn)p;
    }();
    return fn;
}

// Tensor map over a padded four-step workspace as the second pass sees it: [outer * N1 rows][N2 * pitch complex] fp32 pairs,
// box = one tile ([G columns][N1 rows]).  false: fall back to element copies.
bool encode_cols_tensor_map(CUtensorMap* tm, const KParams& k, const void* ws)
{
    encode_tiled_fn enc = tensor_map_encoder();
    if (!enc || ((uintptr_t)ws & 15)) return false;
    const cuuint64_t gdim[2] = {(cuuint64_t)k.inner * 2, (cuuint64_t)k.outer * (cuuint64_t)k.H};
    const cuuint64_t gstride[1] = {(cuuint64_t)k.inner * 8};
    const cuuint32_t box[2] = {(cuuint32_t)k.G * 2, (cuuint32_t)k.H};
    const cuuint32_t estr[2] = {1, 1};
    if ((gstride[0] & 15) || box[0] > 256 || box[1] > 256 || gdim[0] >= (1ull << 32) || gdim[1] >= (1ull << 32)) return false;
    return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(ws), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

struct ResolvedLaunch {
    KParams kp;
    kernel_fn fn;
    int grid;
    int src, dst;
    spec_fn spec = nullptr;          // non-null: launch the specialised kernel instead of fn
    const float2* spec_tab = nullptr;
    int spec_smem = 0, spec_nt = 0, spec_grid = 0;
    bool spec_needs_align16 = false; // pipelined kernels move whole images with TMA bulk copies only
    pipe_fn gpipe = nullptr;         // non-null: persistent pipelined generic kernel instead of fn
    pipe_fn gpipe_pack = nullptr;    // what runs in its place once trt_dft_plan_set_column_split has been called
                                     // (the two-step row programs write plain rows only)
    int gp_ns = 0, gp_ng = 0, gp_stage = 0, gp_smem = 0, gp_grid = 0, gp_tables = 0;
    int gp_threads = dft::kPipeThreads, gp_threads_pack = dft::kPipeThreads;   // CTA size of gpipe / gpipe_pack
    int gp_smem_nopack = 0;          // gp_smem without the staging rows of the bulk-store epilogue
    int gp_ns_pack = 0, gp_ng_pack = 0;   // ring depth / groups while the bulk-store epilogue is on
};

}  // namespace

struct trt_dft_plan_s {
    dft::PlanDesc pd;
    int device = 0;
    bool c2c = false;                     // trt_dft_plan_create_c2c: complex on both sides
    void* pool = nullptr;                 // one device allocation holding every table
    std::vector<ResolvedLaunch> launches;
    // truncated-mode plan (trt_dft_plan_create_truncated): kept block [2*kh][kw] on the complex side
    bool trunc = false;
    int kh = 0, kw = 0, tH = 0, tW = 0;
    dft::trunc_fn t_fwd = nullptr;    // fused kernels that move only the kept modes ...
    dft::trunc_fn t_inv = nullptr;
    int t_threads = 0, t_smem = 0, t_grid = 0;
    float t_scale = 1.f;
    trt_dft_plan_t t_full = nullptr;      // ... or the full transform through the workspace plus a crop / pad kernel
    size_t t_full_ws_off = 0;             // where the full plan's own workspace starts inside ours
    // typed-I/O plan (trt_dft_plan_create_typed): 16-bit data on both sides, fp32 arithmetic.  Either one kernel that
    // converts as it loads and stores (io_fn; t_threads / t_smem / t_grid / t_scale as above), or the fp32 plan t_full
    // between two conversion passes through the workspace: [fp32 input | fp32 output | t_full's workspace].
    int io_type = TRT_DFT_IO_F32;
    spec_fn io_fn = nullptr;
    size_t io_out_off = 0;
    // trt_dft_exec_host scratch (lazily allocated, reused across calls)
    std::mutex host_mu;
    void* h_in = nullptr;
    void* h_out = nullptr;
    void* h_ws = nullptr;
    int64_t h_chunk = 0;                  // images per pipeline chunk
    trt_dft_plan_t h_sub = nullptr;       // plan for one chunk (null: the whole batch is one chunk)
    trt_dft_plan_t h_sub_tail = nullptr;  // plan for the last, shorter chunk
    cudaStream_t h_streams[3] = {nullptr, nullptr, nullptr};
    cudaEvent_t h_events[64] = {};        // two per pipeline chunk (at most 32 chunks)
};

extern "C" {

const char* trt_dft_version(void)
{
    return "b200-dft 0.1 (sm_100a; hand-written mixed-radix shared-memory FFT, no cuFFT)";
}

const char* trt_dft_error_string(int status)
{
    switch (status) {
        case TRT_DFT_OK: return "ok";
        case TRT_DFT_ERR_INVALID_ARG: return "invalid argument";
        case TRT_DFT_ERR_CUDA: return "CUDA error";
        case TRT_DFT_ERR_ALLOC: return "allocation failed";
        case TRT_DFT_ERR_WORKSPACE: return "workspace missing or misaligned";
        case TRT_DFT_ERR_UNSUPPORTED: return "unsupported size";
        default: return "unknown status";
    }
}

static int realize_plan(trt_dft_plan_s* pl, trt_dft_plan_t* plan_out);

int trt_dft_plan_create(trt_dft_plan_t* plan_out, int direction, int rank, const int64_t* dims, int64_t batch)
{
    if (!plan_out) return TRT_DFT_ERR_INVALID_ARG;
    *plan_out = nullptr;
    trt_dft_plan_s* pl = new (std::nothrow) trt_dft_plan_s();
    if (!pl) return TRT_DFT_ERR_ALLOC;
    int rc = TRT_DFT_OK;
    try {
        rc = dft::build_plan(&pl->pd, direction, rank, dims, batch) ? TRT_DFT_ERR_INVALID_ARG : TRT_DFT_OK;
    } catch (...) {
        rc = TRT_DFT_ERR_ALLOC;
    }
    if (rc != TRT_DFT_OK) {
        delete pl;
        return rc;
    }
    return realize_plan(pl, plan_out);
}

int trt_dft_plan_create_c2c(trt_dft_plan_t* plan_out, int direction, int64_t n, int64_t outer, int64_t inner, float scale)
{
    if (!plan_out) return TRT_DFT_ERR_INVALID_ARG;
    *plan_out = nullptr;
    trt_dft_plan_s* pl = new (std::nothrow) trt_dft_plan_s();
    if (!pl) return TRT_DFT_ERR_ALLOC;
    int rc = TRT_DFT_OK;
    try {
        rc = dft::build_c2c_plan(&pl->pd, direction, n, outer, inner, scale) ? TRT_DFT_ERR_INVALID_ARG : TRT_DFT_OK;
    } catch (...) {
        rc = TRT_DFT_ERR_ALLOC;
    }
    if (rc != TRT_DFT_OK) {
        delete pl;
        return rc;
    }
    pl->c2c = true;
    return realize_plan(pl, plan_out);
}

int trt_dft_plan_create_truncated(trt_dft_plan_t* plan_out, int direction, const int64_t* dims, int64_t batch, int64_t keep_h, int64_t keep_w)
{
    if (!plan_out) return TRT_DFT_ERR_INVALID_ARG;
    *plan_out = nullptr;
    if (!dims || (direction != 0 && direction != 1) || batch < 0 || dims[0] < 1 || dims[1] < 1 || dims[0] > (1LL << 30) || dims[1] > (1LL << 30))
        return TRT_DFT_ERR_INVALID_ARG;
    const int64_t H = dims[0], W = dims[1], MO = W / 2 + 1;
    if (keep_h < 1 || 2 * keep_h > H || keep_w < 1 || keep_w > MO) return TRT_DFT_ERR_INVALID_ARG;
    trt_dft_plan_s* pl = new (std::nothrow) trt_dft_plan_s();
    if (!pl) return TRT_DFT_ERR_ALLOC;
    pl->trunc = true;
    pl->kh = (int)keep_h;
    pl->kw = (int)keep_w;
    pl->tH = (int)H;
    pl->tW = (int)W;
    pl->pd.direction = direction;
    pl->pd.rank = 2;
    pl->pd.dims[0] = H;
    pl->pd.dims[1] = W;
    pl->pd.batch = batch;
    const int64_t real_floats = batch * H * W, kept_floats = batch * 2 * keep_h * keep_w * 2;
    pl->pd.in_floats = direction == TRT_DFT_INVERSE ? kept_floats : real_floats;
    pl->pd.out_floats = direction == TRT_DFT_INVERSE ? real_floats : kept_floats;
    pl->t_scale = (float)(1.0 / ((double)H * (double)W));
    int num_sms = 0;
    if (cudaGetDevice(&pl->device) != cudaSuccess ||
        cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, pl->device) != cudaSuccess) {
        cudaGetLastError();
        delete pl;
        return TRT_DFT_ERR_CUDA;
    }
    char desc[192];
    const SpecEntry* e = (batch > 0 && !getenv("TRT_DFT_NO_SPEC")) ? find_spec((int)H, (int)W) : nullptr;
    // (The fused kernels pay for compacting / expanding the kept block in shared memory, but measured on B200 they beat the
    // full kernels plus a crop / zero-pad pass at every kept fraction: 0.44 vs 0.65 ms per pair even when everything is kept.)
    if (e && e->trunc_fwd && e->trunc_inv && !getenv("TRT_DFT_TRUNC_UNFUSED")) {
        pl->t_fwd = e->trunc_fwd;
        pl->t_inv = e->trunc_inv;
        pl->t_threads = (direction == TRT_DFT_INVERSE && e->NT_inv) ? e->NT_inv : e->NT;
        pl->t_smem = direction == TRT_DFT_INVERSE ? e->smem_i : e->smem_f;
        const void* fn = direction == TRT_DFT_INVERSE ? (const void*)e->trunc_inv : (const void*)e->trunc_fwd;
        int occ = 0;
        if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, pl->t_smem) != cudaSuccess ||
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, pl->t_threads, pl->t_smem) != cudaSuccess || occ < 1) {
            cudaGetLastError();
            delete pl;
            return TRT_DFT_ERR_CUDA;
        }
        // the pass-per-radix fused kernels (128 x 128, 64 x 64) read their twiddles from a small device table
        std::vector<float2> tab = spec_tables(*e, direction == TRT_DFT_INVERSE, (double)pl->t_scale);
        if (cudaMalloc(&pl->pool, tab.size() * sizeof(float2)) != cudaSuccess ||
            cudaMemcpy(pl->pool, tab.data(), tab.size() * sizeof(float2), cudaMemcpyHostToDevice) != cudaSuccess) {
            cudaGetLastError();
            if (pl->pool) cudaFree(pl->pool);
            delete pl;
            return TRT_DFT_ERR_ALLOC;
        }
        pl->t_grid = spec_grid_for(*e, batch, num_sms, occ);
        snprintf(desc, sizeof desc, "truncated_%s<%d,%d> keep=%dx%d images=%lld [fused, kept modes only] grid=%dx%d", direction ? "irfft2" : "rfft2",
                 (int)H, (int)W, 2 * pl->kh, pl->kw, (long long)batch, pl->t_grid, pl->t_threads);
    } else if (batch > 0) {
        const int rc = trt_dft_plan_create(&pl->t_full, direction, 2, dims, batch);
        if (rc != TRT_DFT_OK) {
            delete pl;
            return rc;
        }
        const size_t full_bytes = ((size_t)batch * H * MO * 8 + 255) / 256 * 256;
        pl->t_full_ws_off = full_bytes;
        pl->pd.ws_bytes = full_bytes + trt_dft_workspace_bytes(pl->t_full);
        snprintf(desc, sizeof desc, "truncated_%s<%d,%d> keep=%dx%d images=%lld [full transform through the workspace + %s]",
                 direction ? "irfft2" : "rfft2", (int)H, (int)W, 2 * pl->kh, pl->kw, (long long)batch, direction ? "zero-pad" : "crop");
    } else {
        snprintf(desc, sizeof desc, "empty");
    }
    pl->pd.description = desc;
    *plan_out = pl;
    return TRT_DFT_OK;
}

int trt_dft_plan_create_typed(trt_dft_plan_t* plan_out, int direction, int rank, const int64_t* dims, int64_t batch, int io_type)
{
    if (io_type == TRT_DFT_IO_F32) return trt_dft_plan_create(plan_out, direction, rank, dims, batch);
    if (!plan_out || (io_type != TRT_DFT_IO_F16 && io_type != TRT_DFT_IO_BF16)) return TRT_DFT_ERR_INVALID_ARG;
    *plan_out = nullptr;
    trt_dft_plan_t full = nullptr;
    const int rc = trt_dft_plan_create(&full, direction, rank, dims, batch);     // validates every argument
    if (rc != TRT_DFT_OK) return rc;
    auto* pl = new (std::nothrow) trt_dft_plan_s();
    if (!pl) {
        trt_dft_plan_destroy(full);
        return TRT_DFT_ERR_ALLOC;
    }
    pl->io_type = io_type;
    pl->device = full->device;
    pl->pd.direction = direction;
    pl->pd.rank = rank;
    for (int i = 0; i < rank; ++i) pl->pd.dims[i] = dims[i];
    pl->pd.batch = batch;
    pl->pd.in_floats = full->pd.in_floats;         // element counts; the elements are 16-bit
    pl->pd.out_floats = full->pd.out_floats;
    const char* tname = io_type == TRT_DFT_IO_F16 ? "f16" : "bf16";
    char desc[512];
    const bool inv = direction == TRT_DFT_INVERSE;
    const SpecEntry* e = (rank == 2 && batch > 0 && !getenv("TRT_DFT_NO_SPEC") && !getenv("TRT_DFT_IO_UNFUSED"))
                             ? find_spec((int)dims[0], (int)dims[1]) : nullptr;
    const spec_fn fused = e ? (inv ? e->inv16[io_type - 1] : e->fwd16[io_type - 1]) : nullptr;
    int num_sms = 0, occ = 0;
    if (fused && cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, pl->device) == cudaSuccess) {
        pl->t_threads = (inv && e->NT_inv) ? e->NT_inv : e->NT;
        pl->t_smem = inv ? e->smem_i : e->smem_f;
        pl->t_scale = (float)(1.0 / ((double)dims[0] * (double)dims[1]));
        if (cudaFuncSetAttribute((const void*)fused, cudaFuncAttributeMaxDynamicSharedMemorySize, pl->t_smem) == cudaSuccess &&
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void*)fused, pl->t_threads, pl->t_smem) == cudaSuccess && occ >= 1) {
            // the pass-per-radix fused kernels (128 x 128, 64 x 64) read their twiddles from a small device table
            std::vector<float2> tab = spec_tables(*e, inv, (double)pl->t_scale);
            if (cudaMalloc(&pl->pool, tab.size() * sizeof(float2)) == cudaSuccess &&
                cudaMemcpy(pl->pool, tab.data(), tab.size() * sizeof(float2), cudaMemcpyHostToDevice) == cudaSuccess) {
                pl->io_fn = fused;
                pl->t_grid = spec_grid_for(*e, batch, num_sms, occ);
            } else {
                cudaGetLastError();
                if (pl->pool) cudaFree(pl->pool);
                pl->pool = nullptr;
            }
        } else {
            cudaGetLastError();
        }
    }
    if (pl->io_fn) {
        trt_dft_plan_destroy(full);
        snprintf(desc, sizeof desc, "%s<%d,%d> %s in, %s out, fp32 arithmetic images=%lld [converts as it loads and stores] grid=%dx%d smem=%d",
                 inv ? "irfft2" : "rfft2", (int)dims[0], (int)dims[1], tname, tname, (long long)batch, pl->t_grid, pl->t_threads, pl->t_smem);
    } else {
        pl->t_full = full;
        const size_t a = ((size_t)full->pd.in_floats * 4 + 255) / 256 * 256, b = ((size_t)full->pd.out_floats * 4 + 255) / 256 * 256;
        pl->io_out_off = a;
        pl->t_full_ws_off = a + b;
        pl->pd.ws_bytes = batch > 0 ? a + b + full->pd.ws_bytes : 0;
        snprintf(desc, sizeof desc, "%s -> fp32 ; %.380s ; fp32 -> %s", tname, full->pd.description.c_str(), tname);
    }
    pl->pd.description = desc;
    *plan_out = pl;
    return TRT_DFT_OK;
}

int trt_dft_plan_io_type(trt_dft_plan_t plan) { return plan ? plan->io_type : -1; }

// Device side of plan creation: tables to the GPU, kernel attributes, grid sizes.
static int realize_plan(trt_dft_plan_s* pl, trt_dft_plan_t* plan_out)
{
    if (pl->pd.launches.empty()) {   // batch == 0
        *plan_out = pl;
        return TRT_DFT_OK;
    }
    int num_sms = 0;
    if (cudaGetDevice(&pl->device) != cudaSuccess ||
        cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, pl->device) != cudaSuccess) {
        delete pl;
        return TRT_DFT_ERR_CUDA;
    }
    // shape-specialised fused kernels bring their own (small) table
    std::vector<int> spec_tab_id(pl->pd.launches.size(), -1);
    std::vector<const SpecEntry*> spec_of(pl->pd.launches.size(), nullptr);
    for (size_t i = 0; i < pl->pd.launches.size(); ++i) {
        const KParams& k = pl->pd.launches[i].kp;
        if (k.mode != dft::MODE_F2D && k.mode != dft::MODE_I2D) continue;
        if (getenv("TRT_DFT_NO_SPEC")) continue;     // benchmarking override: force the generic engine
        const SpecEntry* e = find_spec(k.H, k.W);
        if (!e) continue;
        spec_of[i] = e;
        pl->pd.tw_tables.push_back(spec_tables(*e, k.mode == dft::MODE_I2D, (double)k.scale));
        spec_tab_id[i] = (int)pl->pd.tw_tables.size() - 1;
        pl->pd.launches[i].name += " [specialised]";
    }
    {
        std::string d;
        for (size_t i = 0; i < pl->pd.launches.size(); ++i) d += (i ? " ; " : "") + pl->pd.launches[i].name;
        pl->pd.description = d;
    }
    // pack all tables into one allocation
    std::vector<size_t> tw_off(pl->pd.tw_tables.size()), perm_off(pl->pd.perm_tables.size());
    size_t total = 0;
    auto bump = [&](size_t bytes) { size_t o = total; total += (bytes + 255) / 256 * 256; return o; };
    for (size_t i = 0; i < tw_off.size(); ++i) tw_off[i] = bump(pl->pd.tw_tables[i].size() * sizeof(float2));
    for (size_t i = 0; i < perm_off.size(); ++i) perm_off[i] = bump(pl->pd.perm_tables[i].size() * sizeof(int));
    std::vector<unsigned char> staging(total ? total : 256, 0);
    for (size_t i = 0; i < tw_off.size(); ++i)
        memcpy(staging.data() + tw_off[i], pl->pd.tw_tables[i].data(), pl->pd.tw_tables[i].size() * sizeof(float2));
    for (size_t i = 0; i < perm_off.size(); ++i)
        memcpy(staging.data() + perm_off[i], pl->pd.perm_tables[i].data(), pl->pd.perm_tables[i].size() * sizeof(int));
    if (cudaMalloc(&pl->pool, staging.size()) != cudaSuccess) {
        cudaGetLastError();
        delete pl;
        return TRT_DFT_ERR_ALLOC;
    }
    if (cudaMemcpy(pl->pool, staging.data(), staging.size(), cudaMemcpyHostToDevice) != cudaSuccess) {
        cudaFree(pl->pool);
        delete pl;
        return TRT_DFT_ERR_CUDA;
    }
    unsigned char* base = (unsigned char*)pl->pool;
    for (size_t li = 0; li < pl->pd.launches.size(); ++li) {
        auto& l = pl->pd.launches[li];
        ResolvedLaunch r;
        r.kp = l.kp;
        if (l.tw_row_id >= 0) r.kp.tw_row = (const float2*)(base + tw_off[l.tw_row_id]);
        if (l.tw_row_t_id >= 0) r.kp.tw_row_t = (const float2*)(base + tw_off[l.tw_row_t_id]);
        if (l.tw_col_id >= 0) r.kp.tw_col = (const float2*)(base + tw_off[l.tw_col_id]);
        if (l.tw_unt_id >= 0) r.kp.tw_unt = (const float2*)(base + tw_off[l.tw_unt_id]);
        if (l.perm_row_id >= 0) r.kp.perm_row = (const int*)(base + perm_off[l.perm_row_id]);
        if (l.perm_col_id >= 0) r.kp.perm_col = (const int*)(base + perm_off[l.perm_col_id]);
        if (l.tw_four_id >= 0) r.kp.tw_four = (const float2*)(base + tw_off[l.tw_four_id]);
        const bool small_rows = r.kp.mode == dft::MODE_SMALL_ROWS_R2C || r.kp.mode == dft::MODE_SMALL_ROWS_C2R;
        if (r.kp.mode != dft::MODE_C2C_COLS && !small_rows) r.kp.threads = 256;
        r.fn = small_rows ? dft::small_rows_kernel_for(r.kp.mode, r.kp.M) : dft::tile_kernel_for(r.kp.mode, r.kp.threads);
        r.src = l.src;
        r.dst = l.dst;
        if (!r.fn) {
            cudaFree(pl->pool);
            delete pl;
            return TRT_DFT_ERR_UNSUPPORTED;
        }
        if (spec_of[li]) {
            const SpecEntry* e = spec_of[li];
            const bool invm = r.kp.mode == dft::MODE_I2D;
            r.spec = invm ? e->inv : e->fwd;
            r.spec_smem = invm ? e->smem_i : e->smem_f;
            r.spec_nt = (invm && e->NT_inv) ? e->NT_inv : e->NT;
            r.spec_tab = (const float2*)(base + tw_off[spec_tab_id[li]]);
            int occ = 0;
            if (cudaFuncSetAttribute((const void*)r.spec, cudaFuncAttributeMaxDynamicSharedMemorySize, r.spec_smem) != cudaSuccess ||
                cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void*)r.spec, r.spec_nt, r.spec_smem) != cudaSuccess ||
                occ < 1) {
                cudaGetLastError();
                cudaFree(pl->pool);
                delete pl;
                return TRT_DFT_ERR_CUDA;
            }
            r.spec_grid = spec_grid_for(*e, r.kp.nunits, num_sms, occ);
            r.spec_needs_align16 = e->pipelined;
        }
        // the generic engine launch is always prepared: it is the route for buffers the
        // specialised kernels cannot take (not 16-byte aligned)
        if (small_rows) {
            r.grid = (int)r.kp.ntiles;
        } else if (r.kp.mode >= dft::MODE_NAIVE_ROWS_R2C) {
            r.kp.smem_bytes = 0;
            const long long blocks = (r.kp.nunits + 255) / 256;
            r.grid = (int)std::min<long long>(blocks, (long long)num_sms * 16);
        } else {
            if (r.kp.smem_bytes > 48 * 1024 &&
                cudaFuncSetAttribute((const void*)r.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, dft::kSmemMax) !=
                    cudaSuccess) {
                cudaGetLastError();
                cudaFree(pl->pool);
                delete pl;
                return TRT_DFT_ERR_CUDA;
            }
            int occ = 0;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void*)r.fn, r.kp.threads, r.kp.smem_bytes) !=
                    cudaSuccess || occ < 1) {
                cudaGetLastError();
                cudaFree(pl->pool);
                delete pl;
                return TRT_DFT_ERR_CUDA;
            }
            // (same rule as spec_grid_for: beyond two full rounds one CTA per tile beats the fixed round-robin -- 2048 images of
            // 60 x 120 0.066 -> 0.062 ms, 8192 of 32 x 32 0.066 -> 0.055, 4096 of 34 x 74 0.214 -> 0.182)
            {
                const long long slots = (long long)num_sms * occ;
                const bool per_tile = r.kp.ntiles > 2 * slots && r.kp.ntiles <= 0x7fffffffLL && !getenv("TRT_DFT_SPEC_PERSISTENT");
                r.grid = per_tile ? (int)r.kp.ntiles : (int)std::min<long long>(r.kp.ntiles, slots);
            }
            // persistent pipelined variant for the big row / strided-axis passes
            pipe_fn pf = dft::pipe_kernel_for(r.kp.mode);
            const int stage = (r.kp.smem_bytes - dft::kBarrierBytes + 127) / 128 * 128;
            const int ns_cap = getenv("TRT_DFT_GP_NS") ? atoi(getenv("TRT_DFT_GP_NS")) : 4;   // tuning override
            int ns = std::min(ns_cap, (dft::kSmemMax - 256) / std::max(stage, 1));
            // tables (twiddles, permutation) ride in shared memory when they fit beside >= 2 stages
            const bool m2d = r.kp.mode == dft::MODE_F2D || r.kp.mode == dft::MODE_I2D;
            const int tab_bytes = r.kp.mode == dft::MODE_C2C_COLS
                                      ? r.kp.col.n * 12 + 64
                                      : r.kp.row.n * 12 + (r.kp.M + 1) * 8 + 64 + (m2d ? r.kp.col.n * 12 + 64 : 0);
            int tables = 0;
            if (ns >= 2) {
                int ns_t = std::min(ns_cap, (dft::kSmemMax - 256 - tab_bytes) / std::max(stage, 1));
                if (ns_t >= 2 && (ns_t >= 3 || ns < 3)) {
                    ns = ns_t;
                    tables = 1;
                }
            }
            int ct = dft::kPipeCompute, ptn = dft::kPipeProducers;
            if (pf && !getenv("TRT_DFT_NO_GEO")) {
                pipe_fn pg = dft::pipe_kernel_for_geo(r.kp, tables != 0, &ct, &ptn);
                if (pg) pf = pg;
                else ct = dft::kPipeCompute, ptn = dft::kPipeProducers;
            }
            if (pf && !getenv("TRT_DFT_NO_GPIPE") && stage >= 8 * 1024 && ns >= 2 && r.kp.ntiles >= num_sms) {
                const int smem = ns * stage + 2 * ns * 8 + 64 + (tables ? tab_bytes : 0);
                r.gp_tables = tables;
                if (cudaFuncSetAttribute((const void*)pf, cudaFuncAttributeMaxDynamicSharedMemorySize, dft::kSmemMax) == cudaSuccess) {
                    r.gpipe = pf;
                    r.gpipe_pack = pf;
                    r.gp_threads = r.gp_threads_pack = ct + ptn;
                    if (r.kp.mode == dft::MODE_F1D) {
                        KParams kpack = r.kp;
                        kpack.pack_n = 1;
                        pipe_fn pp = dft::pipe_kernel_for(r.kp.mode);
                        int ct2 = dft::kPipeCompute, pt2 = dft::kPipeProducers;
                        if (!getenv("TRT_DFT_NO_GEO")) {
                            if (pipe_fn pg2 = dft::pipe_kernel_for_geo(kpack, tables != 0, &ct2, &pt2)) pp = pg2;
                            else ct2 = dft::kPipeCompute, pt2 = dft::kPipeProducers;
                        }
                        r.gp_threads_pack = ct2 + pt2;
                        if (pp != pf && cudaFuncSetAttribute((const void*)pp, cudaFuncAttributeMaxDynamicSharedMemorySize, dft::kSmemMax) != cudaSuccess) {
                            cudaGetLastError();
                            pp = nullptr;            // fall back to the non-persistent kernel for split output
                        }
                        r.gpipe_pack = pp;
                    }
                    r.gp_ns = ns;
                    r.gp_ng = ns >= 3 ? 2 : 1;
                    if (getenv("TRT_DFT_GP_NG")) {
                        // A/B switch.  Groups are whole warps of equal count (named barriers count in warps).
                        const int want = std::max(1, std::min(atoi(getenv("TRT_DFT_GP_NG")), ns - 1));
                        if ((ct / 32) % want == 0) r.gp_ng = want;
                    }
                    r.gp_stage = stage;
                    r.gp_smem = smem;
                    r.gp_smem_nopack = smem;
                    r.gp_grid = (int)std::min<long long>(r.kp.ntiles, num_sms);
                } else {
                    cudaGetLastError();
                }
            }
        }
        if (r.grid < 1) r.grid = 1;
        pl->launches.push_back(r);
    }
    *plan_out = pl;
    return TRT_DFT_OK;
}

// A plan for a few short rows runs one thread per row and only writes plain rows: put it back on the tile engine.
static int replan_on_the_engine(trt_dft_plan_t plan)
{
    trt_dft_plan_s* np = new (std::nothrow) trt_dft_plan_s();
    if (!np) return TRT_DFT_ERR_ALLOC;
    try {
        if (dft::build_plan(&np->pd, plan->pd.direction, plan->pd.rank, plan->pd.dims, plan->pd.batch, dft::PLAN_NO_SMALL_ROWS)) {
            delete np;
            return TRT_DFT_ERR_INVALID_ARG;
        }
    } catch (...) {
        delete np;
        return TRT_DFT_ERR_ALLOC;
    }
    trt_dft_plan_t realized = nullptr;
    const int rc = realize_plan(np, &realized);       // (frees np itself on failure)
    if (rc != TRT_DFT_OK) return rc;
    std::swap(plan->pd, realized->pd);
    std::swap(plan->pool, realized->pool);
    std::swap(plan->launches, realized->launches);
    trt_dft_plan_destroy(realized);
    return TRT_DFT_OK;
}

int trt_dft_plan_set_column_split(trt_dft_plan_t plan, int nparts, const int64_t* col_starts, int64_t rows_per_image)
{
    if (!plan || nparts < 0 || nparts > 8 || (nparts > 0 && (!col_starts || rows_per_image < 1))) return TRT_DFT_ERR_INVALID_ARG;
    if (nparts > 0 && plan->launches.size() == 1 && plan->launches[0].kp.mode == dft::MODE_SMALL_ROWS_R2C) {
        const int rc = replan_on_the_engine(plan);
        if (rc != TRT_DFT_OK) return rc;
    }
    if (plan->launches.size() != 1 || plan->launches[0].kp.mode != dft::MODE_F1D) return TRT_DFT_ERR_UNSUPPORTED;
    KParams& k = plan->launches[0].kp;
    const int Mo = k.M + 1;
    if (nparts > 0) {
        if (col_starts[0] != 0 || col_starts[nparts] != Mo || k.nunits % rows_per_image != 0 || k.nunits > (1LL << 30))
            return TRT_DFT_ERR_INVALID_ARG;
        for (int j = 0; j < nparts; ++j)
            if (col_starts[j + 1] < col_starts[j]) return TRT_DFT_ERR_INVALID_ARG;
    }
    k.pack_n = nparts;
    for (int j = 0; nparts > 0 && j <= nparts; ++j) k.pack_c0[j] = (int)col_starts[j];   // (nparts == 0: back to plain rows, col_starts may be null)
    k.pack_hn = (int)rows_per_image;
    k.pack_peer_on = 0;
    k.pack_bulk = 0;
    for (int j = 0; j < 8; ++j) k.pack_pitch[j] = j < nparts ? (int)(col_starts[j + 1] - col_starts[j]) : 0;
    plan->launches[0].gp_smem = plan->launches[0].gp_smem_nopack;
    return TRT_DFT_OK;
}

int trt_dft_plan_set_peer_column_split(trt_dft_plan_t plan, int nparts, const int64_t* col_starts, int64_t rows_per_image,
                                       int64_t image_rows, int64_t row_offset, void* const* peer_slabs, const int64_t* slab_pitches)
{
    if (nparts < 1 || !peer_slabs || image_rows < 1 || row_offset < 0 || row_offset + rows_per_image > image_rows ||
        image_rows > (1LL << 30))
        return TRT_DFT_ERR_INVALID_ARG;
    for (int j = 0; j < nparts && j < 8; ++j)
        if (!peer_slabs[j] || ((uintptr_t)peer_slabs[j] & 7)) return TRT_DFT_ERR_INVALID_ARG;
    const int rc = trt_dft_plan_set_column_split(plan, nparts, col_starts, rows_per_image);
    if (rc != TRT_DFT_OK) return rc;
    ResolvedLaunch& rl = plan->launches[0];
    KParams& k = rl.kp;
    for (int j = 0; j < nparts; ++j) {
        const int64_t cn = col_starts[j + 1] - col_starts[j];
        const int64_t pitch = slab_pitches ? slab_pitches[j] : cn;
        if (pitch < cn || pitch > (1LL << 30)) return TRT_DFT_ERR_INVALID_ARG;
        k.pack_pitch[j] = (int)pitch;
    }
    k.pack_peer_on = 1;
    k.pack_H = (int)image_rows;
    k.pack_h0 = (int)row_offset;
    for (int j = 0; j < 8; ++j) k.pack_peer[j] = j < nparts ? (float2*)peer_slabs[j] : nullptr;
    // Bulk stores (KParams::pack_bulk): every run [c0, c0 + cn) of a staged row must start and end on 16 bytes on both
    // sides -- even column starts, even pitches that cover an odd cn rounded up, 16-byte aligned slabs -- and the staging rows
    // of the compute groups must fit beside the tile ring.
    k.pack_bulk = 0;
    bool ok = rl.gpipe_pack != nullptr && !getenv("TRT_DFT_NO_PEER_BULK");
    for (int j = 0; ok && j < nparts; ++j) {
        const int64_t cn = col_starts[j + 1] - col_starts[j];
        ok = (col_starts[j] % 2 == 0) && (k.pack_pitch[j] % 2 == 0) && (k.pack_pitch[j] >= (cn + 1) / 2 * 2) && (((uintptr_t)peer_slabs[j] & 15) == 0);
    }
    rl.gp_ns_pack = rl.gp_ns;
    rl.gp_ng_pack = rl.gp_ng;
    if (ok) {
        // Room for the staging rows: the split twiddles leave shared memory (the kernel reads them from global memory in
        // this mode) and the tile ring gives up stages until everything fits.
        const int mo_pad = (k.M + 1 + 1) / 2 * 2;                       // staging row: M + 1 results + padding to an even count
        const int stage_bytes = (k.G * mo_pad * 8 + 127) / 128 * 128;
        const int tabs = rl.gp_tables ? k.row.n * 12 + 128 : 0;         // per-step twiddles ((n - 1) * 8) + permutation (n * 4), 16-byte rounding
        for (int ns = rl.gp_ns; ns >= 2 && !k.pack_bulk; --ns) {
            const int ng = std::min(rl.gp_ng, ns >= 3 ? 2 : 1);
            const int base = (ns * rl.gp_stage + 2 * ns * 8 + 64 + tabs + 127) / 128 * 128;
            if (base + ng * stage_bytes > dft::kSmemMax) continue;
            k.pack_bulk = 1;
            k.pack_mo_pad = mo_pad;
            k.pack_stage_off = base;
            k.pack_stage_bytes = stage_bytes;
            rl.gp_ns_pack = ns;
            rl.gp_ng_pack = ng;
            rl.gp_smem = base + ng * stage_bytes;
        }
    }
    if (!k.pack_bulk) rl.gp_smem = rl.gp_smem_nopack;
    return TRT_DFT_OK;
}

int trt_dft_plan_set_peer_column_gather(trt_dft_plan_t plan, int nparts, const int64_t* col_starts, int64_t rows_per_image,
                                        int64_t image_rows, int64_t row_offset, void* const* peer_slabs, const int64_t* slab_pitches)
{
    if (!plan || nparts < 0 || nparts > 8) return TRT_DFT_ERR_INVALID_ARG;
    if (plan->launches.size() != 1 || plan->launches[0].kp.mode != dft::MODE_I1D) return TRT_DFT_ERR_UNSUPPORTED;
    ResolvedLaunch& rl = plan->launches[0];
    KParams& k = rl.kp;
    if (nparts == 0) {
        k.pack_n = 0;
        k.pack_peer_on = 0;
        return TRT_DFT_OK;
    }
    if (!rl.gpipe) return TRT_DFT_ERR_UNSUPPORTED;              // only the persistent row kernel's producers can gather
    const int Mo = k.M + 1;
    if (!col_starts || !peer_slabs || rows_per_image < 1 || image_rows < 1 || row_offset < 0 || row_offset + rows_per_image > image_rows ||
        image_rows > (1LL << 30) || col_starts[0] != 0 || col_starts[nparts] != Mo || k.nunits % rows_per_image != 0 || k.nunits > (1LL << 30))
        return TRT_DFT_ERR_INVALID_ARG;
    for (int j = 0; j < nparts; ++j) {
        const int64_t cn = col_starts[j + 1] - col_starts[j];
        const int64_t pitch = slab_pitches ? slab_pitches[j] : cn;
        if (cn < 0 || pitch < cn || pitch > (1LL << 30) || !peer_slabs[j] || ((uintptr_t)peer_slabs[j] & 7)) return TRT_DFT_ERR_INVALID_ARG;
        k.pack_pitch[j] = (int)pitch;
    }
    k.pack_n = nparts;
    for (int j = 0; j <= nparts; ++j) k.pack_c0[j] = (int)col_starts[j];
    k.pack_hn = (int)rows_per_image;
    k.pack_H = (int)image_rows;
    k.pack_h0 = (int)row_offset;
    k.pack_peer_on = 1;
    k.pack_bulk = 0;
    for (int j = 0; j < 8; ++j) k.pack_peer[j] = j < nparts ? (float2*)peer_slabs[j] : nullptr;
    return TRT_DFT_OK;
}

int trt_dft_plan_set_peer_row_split(trt_dft_plan_t plan, int nparts, const int64_t* row_starts, int64_t dest_pitch,
                                    int64_t dest_col_offset, void* const* peer_bands)
{
    if (!plan || !plan->c2c || nparts < 0 || nparts > 8) return TRT_DFT_ERR_INVALID_ARG;
    if (plan->launches.empty()) return TRT_DFT_OK;
    ResolvedLaunch& last = plan->launches.back();                 // the pass that writes the result
    if (last.kp.mode != dft::MODE_C2C_COLS || last.kp.n1split > 1) return TRT_DFT_ERR_UNSUPPORTED;
    KParams& k = last.kp;
    if (nparts == 0) {
        k.split_n = 0;
        return TRT_DFT_OK;
    }
    // second pass of a four-step FFT: rows and columns of the image are folded differently (KParams::split_n2)
    const bool four = plan->launches.size() == 2 && plan->launches[0].kp.n1split > 1;
    const long long n2 = four ? plan->launches[0].kp.H : 1;
    const long long n_axis = (long long)k.H * n2, inner = k.inner / n2;
    const long long width = k.unflat_w ? k.unflat_w : inner;     // (a padded workspace's extra column is never stored)
    if (!row_starts || !peer_bands || row_starts[0] != 0 || row_starts[nparts] != n_axis || dest_pitch < 1 || dest_col_offset < 0 ||
        dest_col_offset + width > dest_pitch || dest_pitch > (1LL << 30))
        return TRT_DFT_ERR_INVALID_ARG;
    k.split_n2 = (int)n2;
    k.split_inner = (int)inner;
    for (int j = 0; j < nparts; ++j)
        if (row_starts[j + 1] < row_starts[j] || !peer_bands[j] || ((uintptr_t)peer_bands[j] & 7)) return TRT_DFT_ERR_INVALID_ARG;
    k.split_n = nparts;
    for (int j = 0; j <= nparts; ++j) k.split_r0[j] = (int)row_starts[j];
    k.split_pitch = (int)dest_pitch;
    k.split_c0 = (int)dest_col_offset;
    for (int j = 0; j < 8; ++j) k.split_peer[j] = j < nparts ? (float2*)peer_bands[j] : nullptr;
    return TRT_DFT_OK;
}

size_t trt_dft_workspace_bytes(trt_dft_plan_t plan)
{
    return plan ? plan->pd.ws_bytes : 0;
}

size_t trt_dft_query_workspace_bytes(int direction, int rank, const int64_t* dims, int64_t batch)
{
    try {
        dft::PlanDesc pd;
        if (dft::build_plan(&pd, direction, rank, dims, batch)) return 0;
        return pd.ws_bytes;
    } catch (...) {
        return 0;
    }
}

int64_t trt_dft_plan_input_floats(trt_dft_plan_t plan) { return plan ? plan->pd.in_floats : -1; }
int64_t trt_dft_plan_output_floats(trt_dft_plan_t plan) { return plan ? plan->pd.out_floats : -1; }

int trt_dft_plan_describe(trt_dft_plan_t plan, char* buf, size_t len)
{
    if (!plan) return -1;
    if (buf && len) {
        std::string d = plan->pd.description;
        for (size_t i = 0; i < plan->launches.size(); ++i) {
            char t[128];
            const auto& rl = plan->launches[i];
            if (rl.gpipe && !rl.spec)
                snprintf(t, sizeof t, "%s[%d]pipelined grid=%dx%d stages=%d groups=%d%s", i ? "," : " ", (int)i, rl.gp_grid,
                         rl.gp_threads, rl.kp.pack_bulk ? rl.gp_ns_pack : rl.gp_ns, rl.kp.pack_bulk ? rl.gp_ng_pack : rl.gp_ng,
                         rl.kp.pack_bulk ? " bulk-stores-to-peers" : (rl.kp.pack_peer_on ? " stores-to-peers" : ""));
            else
                snprintf(t, sizeof t, rl.spec ? "%s[%d]grid=%dx%d smem=%d" : "%s[%d]grid=%dx%d", i ? "," : " ", (int)i,
                         rl.spec ? rl.spec_grid : rl.grid, rl.spec ? rl.spec_nt : rl.kp.threads, rl.spec_smem);
            d += t;
        }
        strncpy(buf, d.c_str(), len - 1);
        buf[len - 1] = 0;
    }
    if (plan->trunc) return plan->pd.batch == 0 ? 0 : (plan->t_fwd ? 1 : 1 + (int)plan->t_full->launches.size());
    if (plan->io_type != TRT_DFT_IO_F32) return plan->pd.batch == 0 ? 0 : (plan->io_fn ? 1 : 2 + (int)plan->t_full->launches.size());
    return (int)plan->launches.size();
}

static int exec_truncated(trt_dft_plan_t plan, const void* in, void* out, void* workspace, void* stream)
{
    if (plan->pd.batch == 0) return TRT_DFT_OK;
    if (!in || !out) return TRT_DFT_ERR_INVALID_ARG;
    int cur_dev = -1;
    if (cudaGetDevice(&cur_dev) != cudaSuccess || cur_dev != plan->device) {
        cudaGetLastError();
        return TRT_DFT_ERR_INVALID_ARG;
    }
    const bool inv = plan->pd.direction == TRT_DFT_INVERSE;
    cudaStream_t st = (cudaStream_t)stream;
    const bool aligned = ((((uintptr_t)in) | ((uintptr_t)out)) & 15) == 0;
    if (plan->t_fwd && aligned) {
        if (!inv)
            plan->t_fwd<<<plan->t_grid, plan->t_threads, plan->t_smem, st>>>((const float*)in, (float*)out, plan->pd.batch,
                                                                            (const float2*)plan->pool, 1.0f, plan->kh, plan->kw);
        else
            plan->t_inv<<<plan->t_grid, plan->t_threads, plan->t_smem, st>>>((const float*)in, (float*)out, plan->pd.batch,
                                                                            (const float2*)plan->pool, plan->t_scale, plan->kh, plan->kw);
        if (cudaPeekAtLastError() != cudaSuccess) {
            cudaGetLastError();
            return TRT_DFT_ERR_CUDA;
        }
        return TRT_DFT_OK;
    }
    if (!plan->t_full) return TRT_DFT_ERR_INVALID_ARG;      // fused kernels only, and the buffers are not 16-byte aligned
    if (!workspace || ((uintptr_t)workspace & 15)) return TRT_DFT_ERR_WORKSPACE;
    if (((uintptr_t)(inv ? in : (const void*)out)) & 7) return TRT_DFT_ERR_INVALID_ARG;
    void* full = workspace;
    void* inner_ws = (unsigned char*)workspace + plan->t_full_ws_off;
    const int MO = plan->tW / 2 + 1;
    if (!inv) {
        const int rc = trt_dft_exec(plan->t_full, in, full, inner_ws, stream);
        if (rc != TRT_DFT_OK) return rc;
        dft::launch_crop_modes(full, out, plan->pd.batch, plan->tH, MO, plan->kh, plan->kw, stream);
    } else {
        dft::launch_pad_modes(in, full, plan->pd.batch, plan->tH, MO, plan->kh, plan->kw, stream);
        const int rc = trt_dft_exec(plan->t_full, full, out, inner_ws, stream);
        if (rc != TRT_DFT_OK) return rc;
    }
    if (cudaPeekAtLastError() != cudaSuccess) {
        cudaGetLastError();
        return TRT_DFT_ERR_CUDA;
    }
    return TRT_DFT_OK;
}

static int exec_typed(trt_dft_plan_t plan, const void* in, void* out, void* workspace, void* stream)
{
    if (plan->pd.batch == 0) return TRT_DFT_OK;
    if (!in || !out) return TRT_DFT_ERR_INVALID_ARG;
    int cur_dev = -1;
    if (cudaGetDevice(&cur_dev) != cudaSuccess || cur_dev != plan->device) {
        cudaGetLastError();
        return TRT_DFT_ERR_INVALID_ARG;
    }
    if (plan->io_fn) {
        // whole images move as bulk copies: 16-byte aligned tensors (any tensor of a framework allocator is)
        if ((((uintptr_t)in) | ((uintptr_t)out)) & 15) return TRT_DFT_ERR_INVALID_ARG;
        plan->io_fn<<<plan->t_grid, plan->t_threads, plan->t_smem, (cudaStream_t)stream>>>((const float*)in, (float*)out, plan->pd.batch,
                                                                                           (const float2*)plan->pool, plan->t_scale);
    } else {
        if (!workspace || ((uintptr_t)workspace & 15)) return TRT_DFT_ERR_WORKSPACE;
        if ((((uintptr_t)in) | ((uintptr_t)out)) & 1) return TRT_DFT_ERR_INVALID_ARG;
        float* a = (float*)workspace;
        float* b = (float*)((unsigned char*)workspace + plan->io_out_off);
        dft::launch_widen(in, a, plan->pd.in_floats, plan->io_type, stream);
        const int rc = trt_dft_exec(plan->t_full, a, b, (unsigned char*)workspace + plan->t_full_ws_off, stream);
        if (rc != TRT_DFT_OK) return rc;
        dft::launch_narrow(b, out, plan->pd.out_floats, plan->io_type, stream);
    }
    if (cudaPeekAtLastError() != cudaSuccess) {
        cudaGetLastError();
        return TRT_DFT_ERR_CUDA;
    }
    return TRT_DFT_OK;
}

int trt_dft_exec(trt_dft_plan_t plan, const void* in, void* out, void* workspace, void* stream)
{
    if (!plan) return TRT_DFT_ERR_INVALID_ARG;
    if (plan->trunc) return exec_truncated(plan, in, out, workspace, stream);
    if (plan->io_type != TRT_DFT_IO_F32) return exec_typed(plan, in, out, workspace, stream);
    if (plan->launches.empty()) return TRT_DFT_OK;
    if (!in || !out) return TRT_DFT_ERR_INVALID_ARG;
    // the plan's tables live on the GPU it was created on: launching from another current device would
    // hand that GPU's kernels pointers into this one's memory
    int cur_dev = -1;
    if (cudaGetDevice(&cur_dev) != cudaSuccess) {
        cudaGetLastError();
        return TRT_DFT_ERR_CUDA;
    }
    if (cur_dev != plan->device) return TRT_DFT_ERR_INVALID_ARG;
    // complex tensors need 8-byte alignment (float2 accesses), real tensors 4-byte
    const bool inv = plan->pd.direction == TRT_DFT_INVERSE;
    const uintptr_t cptr = (uintptr_t)(inv ? in : (const void*)out);
    const uintptr_t rptr = (uintptr_t)(inv ? (const void*)out : in);
    if ((cptr & 7) || (rptr & 3) || (plan->c2c && (rptr & 7))) return TRT_DFT_ERR_INVALID_ARG;
    if (plan->pd.ws_bytes && (!workspace || ((uintptr_t)workspace & 15))) return TRT_DFT_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    for (auto& r : plan->launches) {
        auto sel = [&](int s) -> void* {
            switch (s) {
                case dft::BUF_IN: return const_cast<void*>(in);
                case dft::BUF_OUT: return out;
                case dft::BUF_WS: return workspace;
                default: return (unsigned char*)workspace + plan->pd.ws1_offset;
            }
        };
        const float* src = (const float*)sel(r.src);
        float* dst = (float*)sel(r.dst);
        // real rows that are not 8-byte aligned cannot be written/read as float2 pairs
        if ((r.kp.mode == dft::MODE_I1D || r.kp.mode == dft::MODE_I2D || r.kp.mode == dft::MODE_SMALL_ROWS_C2R) && ((uintptr_t)dst & 7))
            return TRT_DFT_ERR_INVALID_ARG;
        if (r.spec && !(r.spec_needs_align16 && ((((uintptr_t)src) | ((uintptr_t)dst)) & 15)))
            r.spec<<<r.spec_grid, r.spec_nt, r.spec_smem, st>>>(src, dst, r.kp.nunits, r.spec_tab, r.kp.scale);
        else if (r.gpipe && (r.kp.pack_n == 0 || r.gpipe_pack)) {
            CUtensorMap tmap;
            memset(&tmap, 0, sizeof tmap);
            KParams kp = r.kp;
            // the second four-step pass fetches its tiles from the padded workspace with one 2-D TMA tensor copy each
            if (kp.tma_cols && !(kp.mode == dft::MODE_C2C_COLS && encode_cols_tensor_map(&tmap, kp, src))) kp.tma_cols = 0;
            (kp.pack_n ? r.gpipe_pack : r.gpipe)<<<r.gp_grid, kp.pack_n ? r.gp_threads_pack : r.gp_threads, r.gp_smem, st>>>(kp, src, dst, kp.pack_bulk ? r.gp_ns_pack : r.gp_ns,
                                                                                          kp.pack_bulk ? r.gp_ng_pack : r.gp_ng, r.gp_stage,
                                                                                          r.gp_tables, tmap);
        } else
            r.fn<<<r.grid, r.kp.threads, r.kp.smem_bytes, st>>>(r.kp, src, dst);
        if (cudaPeekAtLastError() != cudaSuccess) {
            cudaGetLastError();
            return TRT_DFT_ERR_CUDA;
        }
    }
    return TRT_DFT_OK;
}

int trt_dft_plan_destroy(trt_dft_plan_t plan)
{
    if (!plan) return TRT_DFT_OK;
    int rc = TRT_DFT_OK;
    if (plan->t_full) trt_dft_plan_destroy(plan->t_full);
    if (plan->h_sub) trt_dft_plan_destroy(plan->h_sub);
    if (plan->h_sub_tail) trt_dft_plan_destroy(plan->h_sub_tail);
    for (cudaStream_t st : plan->h_streams)
        if (st) cudaStreamDestroy(st);
    for (cudaEvent_t ev : plan->h_events)
        if (ev) cudaEventDestroy(ev);
    for (void* p : {plan->pool, plan->h_in, plan->h_out, plan->h_ws})
        if (p && cudaFree(p) != cudaSuccess) rc = TRT_DFT_ERR_CUDA;
    delete plan;
    return rc;
}

int trt_dft_exec_host(trt_dft_plan_t plan, const void* host_in, void* host_out)
{
    if (!plan) return TRT_DFT_ERR_INVALID_ARG;
    const bool wrapped = plan->trunc || plan->io_type != TRT_DFT_IO_F32;     // plans without launches of their own
    if (plan->launches.empty() && !(wrapped && plan->pd.batch > 0)) return TRT_DFT_OK;
    if (!host_in || !host_out) return TRT_DFT_ERR_INVALID_ARG;
    int cur_dev = -1;
    if (cudaGetDevice(&cur_dev) != cudaSuccess || cur_dev != plan->device) {
        cudaGetLastError();
        return TRT_DFT_ERR_INVALID_ARG;      // device buffers below must live on the plan's GPU
    }
    std::lock_guard<std::mutex> lock(plan->host_mu);
    const size_t esz = plan->io_type == TRT_DFT_IO_F32 ? 4 : 2;
    const size_t in_bytes = (size_t)plan->pd.in_floats * esz, out_bytes = (size_t)plan->pd.out_floats * esz;
    const int64_t batch = plan->pd.batch;
    if (!plan->h_in) {
        // Large transfers are cut into batch chunks and run as a 3-stream pipeline (H2D | kernels | D2H),
        // so the two PCIe directions and the kernels overlap.
        int nchunks = (in_bytes + out_bytes) >= ((size_t)64 << 20) ? 16 : 1;   // measured at cfg3: 4 -> 82, 8 -> 89, 16 -> 91, 32 -> 89 GB/s
        if (const char* e = getenv("TRT_DFT_HOST_CHUNKS")) nchunks = std::max(1, std::min(32, atoi(e)));   // tuning override
        if (nchunks > batch) nchunks = (int)batch;
        // a c2c plan's `batch` (outer * inner) is not a contiguous leading dimension of [outer][n][inner]: one chunk
        if (plan->c2c || wrapped) nchunks = 1;
        plan->h_chunk = (batch + nchunks - 1) / nchunks;
        const int64_t tail = batch % plan->h_chunk;
        bool ok = cudaMalloc(&plan->h_in, in_bytes) == cudaSuccess && cudaMalloc(&plan->h_out, out_bytes) == cudaSuccess;
        if (ok && plan->h_chunk != batch) {
            ok = trt_dft_plan_create(&plan->h_sub, plan->pd.direction, plan->pd.rank, plan->pd.dims, plan->h_chunk) == TRT_DFT_OK;
            if (ok && tail)
                ok = trt_dft_plan_create(&plan->h_sub_tail, plan->pd.direction, plan->pd.rank, plan->pd.dims, tail) == TRT_DFT_OK;
        }
        const size_t ws = plan->h_sub ? trt_dft_workspace_bytes(plan->h_sub) : plan->pd.ws_bytes;
        ok = ok && (!ws || cudaMalloc(&plan->h_ws, ws) == cudaSuccess);
        for (int i = 0; ok && i < 3; ++i) ok = cudaStreamCreateWithFlags(&plan->h_streams[i], cudaStreamNonBlocking) == cudaSuccess;
        for (int i = 0; ok && i < 64; ++i) ok = cudaEventCreateWithFlags(&plan->h_events[i], cudaEventDisableTiming) == cudaSuccess;
        if (!ok) {
            // leave nothing half-built behind: the next call starts the set-up again from scratch
            cudaGetLastError();
            if (plan->h_sub) trt_dft_plan_destroy(plan->h_sub);
            if (plan->h_sub_tail) trt_dft_plan_destroy(plan->h_sub_tail);
            plan->h_sub = plan->h_sub_tail = nullptr;
            for (cudaStream_t& st : plan->h_streams) {
                if (st) cudaStreamDestroy(st);
                st = nullptr;
            }
            for (cudaEvent_t& ev : plan->h_events) {
                if (ev) cudaEventDestroy(ev);
                ev = nullptr;
            }
            for (void** p : {&plan->h_in, &plan->h_out, &plan->h_ws}) {
                if (*p) cudaFree(*p);
                *p = nullptr;
            }
            plan->h_chunk = 0;
            cudaGetLastError();
            return TRT_DFT_ERR_ALLOC;
        }
    }
    cudaStream_t s_in = plan->h_streams[0], s_run = plan->h_streams[1], s_out = plan->h_streams[2];
    const size_t in_unit = in_bytes / (size_t)batch, out_unit = out_bytes / (size_t)batch;
    int ci = 0;
    for (int64_t b0 = 0; b0 < batch; b0 += plan->h_chunk, ++ci) {
        const int64_t nb = (batch - b0) < plan->h_chunk ? (batch - b0) : plan->h_chunk;
        trt_dft_plan_t sub = plan->h_sub ? (nb == plan->h_chunk ? plan->h_sub : plan->h_sub_tail) : plan;
        unsigned char* din = (unsigned char*)plan->h_in + (size_t)b0 * in_unit;
        unsigned char* dout = (unsigned char*)plan->h_out + (size_t)b0 * out_unit;
        if (cudaMemcpyAsync(din, (const unsigned char*)host_in + (size_t)b0 * in_unit, (size_t)nb * in_unit,
                            cudaMemcpyHostToDevice, s_in) != cudaSuccess) return TRT_DFT_ERR_CUDA;
        cudaEventRecord(plan->h_events[2 * ci], s_in);
        cudaStreamWaitEvent(s_run, plan->h_events[2 * ci], 0);
        const int rc = trt_dft_exec(sub, din, dout, plan->h_ws, s_run);
        if (rc != TRT_DFT_OK) return rc;
        cudaEventRecord(plan->h_events[2 * ci + 1], s_run);
        cudaStreamWaitEvent(s_out, plan->h_events[2 * ci + 1], 0);
        if (cudaMemcpyAsync((unsigned char*)host_out + (size_t)b0 * out_unit, dout, (size_t)nb * out_unit,
                            cudaMemcpyDeviceToHost, s_out) != cudaSuccess) return TRT_DFT_ERR_CUDA;
    }
    if (cudaStreamSynchronize(s_out) != cudaSuccess) return TRT_DFT_ERR_CUDA;
    return TRT_DFT_OK;
}

}  // extern "C"
