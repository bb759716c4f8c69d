// tests/emu/dft_emu.cpp -- TEST-ONLY host emulator of the shared-memory FFT engine.
//
// Compiles the very same planner (dft_planner.h) and tile programs (dft_programs.cuh) that the
// CUDA kernels run, with a one-thread host context in place of a CTA, so the radix schedules,
// permutation tables, real-FFT split/merge formulas and all index arithmetic can be checked
// against the oracle in this GPU-less container (tests/test_engine_emu.py).  It is NOT part of
// libtrt_dft_plugins.so, is never loaded by the product, and is not a fallback: trt_dft_exec
// has no CPU route.
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "dft_planner.h"
#include "dft_programs.cuh"

namespace {
struct HostCtx {
    int tid = 0, nthr = 1;
    void sync() {}
    bool bulk_load(void*, const void*, long long) { return false; }
    float2* stage_out() const { return nullptr; }
    void fence_async_shared() const {}
    void bulk_store(void*, const void*, unsigned) const {}
    void commit_bulk_stores() const {}
    void wait_bulk_stores_read() const {}
};
}  // namespace

extern "C" int dft_emu_exec(int direction, int rank, const int64_t* dims, int64_t batch, const float* in,
                            float* out, char* desc, size_t desc_len)
{
    dft::PlanDesc pd;
    setenv("TRT_DFT_NO_SMALL_ROWS", "1", 1);   // the emulator exercises the tile programs; the one-thread-per-row kernels are covered by the -m gpu tests
    const int rc = dft::build_plan(&pd, direction, rank, dims, batch);
    unsetenv("TRT_DFT_NO_SMALL_ROWS");
    if (rc) return rc;
    if (desc && desc_len) {
        strncpy(desc, pd.description.c_str(), desc_len - 1);
        desc[desc_len - 1] = 0;
    }
    std::vector<char> ws(pd.ws_bytes + 16);
    HostCtx ctx;
    for (auto& l : pd.launches) {
        dft::KParams kp = l.kp;
        if (l.tw_row_id >= 0) kp.tw_row = pd.tw_tables[l.tw_row_id].data();
        if (l.tw_row_t_id >= 0) kp.tw_row_t = pd.tw_tables[l.tw_row_t_id].data();
        if (l.tw_col_id >= 0) kp.tw_col = pd.tw_tables[l.tw_col_id].data();
        if (l.tw_unt_id >= 0) kp.tw_unt = pd.tw_tables[l.tw_unt_id].data();
        if (l.perm_row_id >= 0) kp.perm_row = pd.perm_tables[l.perm_row_id].data();
        if (l.perm_col_id >= 0) kp.perm_col = pd.perm_tables[l.perm_col_id].data();
        if (l.tw_four_id >= 0) kp.tw_four = pd.tw_tables[l.tw_four_id].data();
        auto sel = [&](int s) -> float* {
            switch (s) {
                case dft::BUF_IN: return const_cast<float*>(in);
                case dft::BUF_OUT: return out;
                case dft::BUF_WS: return (float*)ws.data();
                default: return (float*)(ws.data() + pd.ws1_offset);
            }
        };
        const float* src = sel(l.src);
        float* dst = sel(l.dst);
        if (kp.mode >= dft::MODE_NAIVE_ROWS_R2C) {
            for (long long i = 0; i < kp.nunits; ++i) {
                if (kp.mode == dft::MODE_NAIVE_ROWS_R2C) dft::naive_rows_r2c_elem(kp, i, src, dst);
                else if (kp.mode == dft::MODE_NAIVE_ROWS_C2R) dft::naive_rows_c2r_elem(kp, i, src, dst);
                else dft::naive_cols_elem(kp, i, src, dst);
            }
        } else {
            std::vector<float2> buf(kp.smem_bytes / sizeof(float2) + 2);
            for (long long t = 0; t < kp.ntiles; ++t) dft::run_tile(kp, t, buf.data(), src, dst, ctx);
        }
    }
    return 0;
}
