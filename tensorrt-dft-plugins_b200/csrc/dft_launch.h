// dft_launch.h -- kernel lookup functions shared between the executor (dft_kernels.cu) and the translation
// units that instantiate the kernels (dft_tile_kernels.cu, dft_pipe_kernels.cu); split so that nvcc compiles
// them in parallel.
#pragma once
#include <cuda.h>            // CUtensorMap (type only; the encoder is fetched through cudaGetDriverEntryPoint)

#include "dft_types.h"

namespace dft {

typedef void (*kernel_fn)(const KParams, const float*, float*);
typedef void (*pipe_fn)(const KParams, const float*, float*, int, int, int, int, const CUtensorMap);

constexpr int kPipeCompute = 512, kPipeProducers = 64, kPipeThreads = kPipeCompute + kPipeProducers;

kernel_fn tile_kernel_for(int mode, int threads);           // nullptr: unknown mode
kernel_fn small_rows_kernel_for(int mode, int M);           // MODE_SMALL_ROWS_*: one thread per row; nullptr: M not instantiated
pipe_fn pipe_kernel_for(int mode);                          // nullptr: mode has no pipelined variant
pipe_fn pipe_kernel_for_geo(const KParams& k, bool tables_in_smem, int* compute_threads, int* producer_threads);   // nullptr: schedule not registered

// dft_trunc.cu: kept modes <-> full one-sided spectrum (fallback of the truncated-mode plans)
void launch_crop_modes(const void* full, void* kept, long long nimg, int H, int MO, int kh, int kw, void* stream);
void launch_pad_modes(const void* kept, void* full, long long nimg, int H, int MO, int kh, int kw, void* stream);


// dft_typed_io.cu: 16-bit <-> fp32 passes either side of the fp32 transform (fallback of the typed-I/O plans);
// io_type is TRT_DFT_IO_F16 (1) or TRT_DFT_IO_BF16 (2)
void launch_widen(const void* src, float* dst, long long n, int io_type, void* stream);
void launch_narrow(const float* src, void* dst, long long n, int io_type, void* stream);

}  // namespace dft
