/*
 * trt_dft.h -- C ABI of the B200-native DFT kernel library inside libtrt_dft_plugins.so.
 *
 * The reference exports no C symbols at all: its API is "dlopen the .so and the two plugin
 * creators register themselves" (/root/reference/src/dft_plugins/dft_plugins.cpp:575-576,
 * /root/reference/src/trt_dft_plugins/__init__.py:26-32).  Its hot path is the pair of
 * library calls made from enqueue.  Each entry point below replaces one of those call sites,
 * so that the plugin shell's enqueue is a few lines over this ABI and so that the path can be
 * driven (ctypes, cgo, JNI, ...) and tested without TensorRT.
 *
 * Conventions: plain pointers and sizes only; every function returns an int status
 * (TRT_DFT_OK == 0) unless noted; nothing throws, nothing synchronises the device unless the
 * comment says so; all device work is enqueued on the caller's stream.
 * Layouts are the plugin's: fp32, C-contiguous ("kLINEAR"), complex = trailing dim of 2
 * (re, im interleaved), one-sided in the last signal dim, leading dims folded into `batch`
 * (splitSignalDims, dft_plugins.cpp:250-266).
 */
#ifndef TRT_DFT_H_
#define TRT_DFT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TRT_DFT_OK 0
#define TRT_DFT_ERR_INVALID_ARG 1   /* bad direction / rank / dims / null pointer            */
#define TRT_DFT_ERR_CUDA 2          /* a CUDA runtime call or kernel launch failed            */
#define TRT_DFT_ERR_ALLOC 3         /* host or device allocation failed                       */
#define TRT_DFT_ERR_WORKSPACE 4     /* plan needs workspace and none (or misaligned) was given */
#define TRT_DFT_ERR_UNSUPPORTED 5   /* size outside what this build handles                   */

#define TRT_DFT_FORWARD 0           /* R2C, unnormalised      (CUFFT_FORWARD, dft_plugins.cpp:346) */
#define TRT_DFT_INVERSE 1           /* C2R, scaled by 1/prod(dims) (CUFFT_INVERSE + scal, :400,:457-468) */

typedef struct trt_dft_plan_s* trt_dft_plan_t;

/* Library version string, e.g. "b200-dft 0.1 (sm_100a)".  Never NULL. */
const char* trt_dft_version(void);

/* Human-readable text for a status code.  Never NULL. */
const char* trt_dft_error_string(int status);

/*
 * Replaces cufftCreate + cufftSetAutoAllocation + cufftXtMakePlanMany
 * (dft_plugins.cpp:161-176, called from configurePlugin :131-178).
 *   direction : TRT_DFT_FORWARD or TRT_DFT_INVERSE
 *   rank      : signal_ndim, 1..3                                   (dft_plugins.cpp:57)
 *   dims      : the LOGICAL REAL signal size, rank entries -- the input's trailing dims for
 *               Rfft (:396), the output's trailing dims for Irfft (:488); any length >= 1
 *   batch     : product of all leading dims (:260-263), >= 0
 * Builds the radix schedule, uploads twiddle / permutation tables to the current device
 * (cudaMalloc + synchronous cudaMemcpy: call it at configure time, not inside a stream
 * capture) and picks the kernels.  The plan is immutable afterwards and may be used from
 * several streams at once.
 */
int trt_dft_plan_create(trt_dft_plan_t* plan, int direction, int rank, const int64_t* dims, int64_t batch);

/*
 * The strided-axis half of a multi-dimensional plan on its own: complex-to-complex FFT of length n
 * along the middle axis of data[outer][n][inner] (interleaved complex fp32), unnormalised, every
 * stored value multiplied by `scale`.  In the reference this work is hidden inside the rank-2/3
 * cufftXtExec (dft_plugins.cpp:171-176, :192); it is exposed because the pencil-decomposed multi-GPU
 * rfft2 runs it after the all-to-all transpose.  direction TRT_DFT_FORWARD uses
 * exp(-i..), TRT_DFT_INVERSE exp(+i..).  `in` may equal `out` in trt_dft_exec (each tile is read
 * completely before it is written; lengths outside the radix set go through the workspace for that).  Query
 * trt_dft_workspace_bytes: it is non-zero for long axes (>= 4096 points, run as a four-step FFT through a
 * workspace copy of the tensor) and for lengths with a prime factor above 8191 (naive kernel, never in place).
 */
int trt_dft_plan_create_c2c(trt_dft_plan_t* plan, int direction, int64_t n, int64_t outer, int64_t inner, float scale);

/*
 * Pencil decomposition helper (no counterpart in the single-GPU reference, dft_plugins.cpp:341 "assuming
 * single GPU"): makes a rank-1 FORWARD plan over `batch` rows write its result split by columns, ready to
 * be the send buffer of an all-to-all.  Rows are taken in images of `rows_per_image`; inside each image the
 * columns [col_starts[j], col_starts[j+1]) of all its rows form one contiguous block (block j goes to rank
 * j), blocks in rank order.  col_starts has nparts+1 entries, col_starts[0] = 0, col_starts[nparts] =
 * dims[0]/2+1; nparts <= 8; nparts = 0 restores the ordinary layout.  Total output size is unchanged.
 * Call before trt_dft_exec, not concurrently with it.
 */
int trt_dft_plan_set_column_split(trt_dft_plan_t plan, int nparts, const int64_t* col_starts, int64_t rows_per_image);

/*
 * Truncated-mode rfft2 / irfft2 for spectral layers (FNO / AFNO: "why FFT plugins for TensorRT",
 * /root/reference/README.md:3) that keep only a corner of low modes.  No counterpart in the reference, whose
 * Rfft / Irfft always move the full one-sided spectrum (README.md:9-14); this is what fusing the layer's mode
 * selection into the transform buys: the discarded modes never touch HBM.
 *   dims = {H, W} (logical real size), kept modes = rows [0, keep_h) and [H - keep_h, H), columns [0, keep_w)
 *   of the one-sided spectrum [H][W/2+1];  1 <= keep_h <= H/2, 1 <= keep_w <= W/2+1.
 *   FORWARD : in  [batch][H][W] real            -> out [batch][2*keep_h][keep_w] complex
 *             == torch.fft.rfft2(x)[..., rows, :keep_w] with rows = cat(0..keep_h-1, H-keep_h..H-1)
 *   INVERSE : in  [batch][2*keep_h][keep_w] complex -> out [batch][H][W] real, scaled by 1/(H*W)
 *             == torch.fft.irfft2 of the spectrum that holds `in` at those modes and zeros elsewhere.
 * Shapes with fused kernels (90 x 180, 128 x 128, 64 x 64, 64 x 128, 128 x 64) run one kernel that moves only the kept block
 * (no workspace; in/out must be 16-byte aligned); every other shape runs the full transform through the workspace
 * plus a crop / zero-pad kernel (query trt_dft_workspace_bytes).  Executed with trt_dft_exec / trt_dft_exec_host.
 */
int trt_dft_plan_create_truncated(trt_dft_plan_t* plan, int direction, const int64_t* dims, int64_t batch, int64_t keep_h, int64_t keep_w);

/*
 * 16-bit tensors at the plan's boundary, fp32 arithmetic inside.  No counterpart in the reference, which accepts
 * fp32 only (supportsFormatCombination, dft_plugins.cpp:93-105), so a half-precision TensorRT engine has to convert
 * either side of the plugin -- two more passes over HBM for a transform that is bound by HBM.
 *   io_type = TRT_DFT_IO_F32  : exactly trt_dft_plan_create
 *             TRT_DFT_IO_F16 / TRT_DFT_IO_BF16 : the real tensor and the complex tensor (re, im interleaved, trailing
 *             dim of 2) both hold 16-bit elements; same shapes, same scaling as trt_dft_plan_create.  Results are the
 *             fp32 transform of the 16-bit input, rounded to nearest even on the way out.  (fp16 overflows at 65504:
 *             an unnormalised forward transform of n points grows like sqrt(n) times the input's magnitude.)
 * 90 x 180, 128 x 128, 64 x 64, 64 x 128 and 128 x 64 run kernels that convert as they load and store, so only 16-bit data crosses HBM
 * (no workspace; both tensors 16-byte aligned).  Every other shape runs the fp32 plan between two conversion passes through the workspace
 * (query trt_dft_workspace_bytes; tensors 2-byte aligned).  trt_dft_plan_input_floats / _output_floats count elements.
 * Executed with trt_dft_exec / trt_dft_exec_host / trt_dft_exec_static_ws.
 */
#define TRT_DFT_IO_F32 0
#define TRT_DFT_IO_F16 1
#define TRT_DFT_IO_BF16 2
int trt_dft_plan_create_typed(trt_dft_plan_t* plan, int direction, int rank, const int64_t* dims, int64_t batch, int io_type);
/* The io_type a plan was created with (TRT_DFT_IO_F32 for every other constructor), -1 for NULL. */
int trt_dft_plan_io_type(trt_dft_plan_t plan);

/*
 * Multi-GPU pencil rfft2 with the exchange fused into the row pass (no counterpart in the single-GPU reference,
 * dft_plugins.cpp:341).  Same column split as trt_dft_plan_set_column_split, but block j of every image is not
 * packed into `out`: the row kernel stores it straight into rank j's column slab
 *     peer_slabs[j][image][image_rows][col_starts[j+1] - col_starts[j]]   (interleaved complex fp32)
 * at rows row_offset .. row_offset + rows_per_image - 1 -- stores through a peer mapping of that GPU's memory (CUDA
 * IPC / symmetric memory; peer_slabs[own rank] is simply local memory).  `out` of trt_dft_exec is then unused (pass
 * any valid pointer).  The caller synchronises the ranks (all rows written before any rank starts its column pass,
 * and nobody overwrites a slab that is still being read).  nparts <= 8.
 *   slab_pitches : row pitch of each slab in complex elements (NULL: dense, pitch = its column count).
 * When every col_starts[j] is even, every pitch is even and at least the slab's column count rounded up to even, and
 * the slabs are 16-byte aligned, each (row, destination) run leaves shared memory as ONE bulk copy (TMA store into the
 * peer's memory) that overlaps the next rows' transform; an odd column count then also writes the slab's padding
 * column.  Otherwise the kernel falls back to 8-byte stores from its registers.
 */
int trt_dft_plan_set_peer_column_split(trt_dft_plan_t plan, int nparts, const int64_t* col_starts, int64_t rows_per_image,
                                       int64_t image_rows, int64_t row_offset, void* const* peer_slabs, const int64_t* slab_pitches);

/*
 * The pencil irfft2 with the exchange fused into the row pass's LOADS: a rank-1 INVERSE plan over this rank's row band
 * takes row R of image b not from `in` (then unused) but from the ranks' column slabs -- columns
 * [col_starts[j], col_starts[j+1]) of that row from peer_slabs[j][b][row_offset + R][.] (pitch slab_pitches[j]; NULL:
 * dense) -- with the persistent row kernel's producer warps copying straight out of the owners' memory (plain loads over
 * NVLink into shared memory, no multicast).  Same arguments as trt_dft_plan_set_peer_column_split; nparts = 0 switches
 * it off.  TRT_DFT_ERR_UNSUPPORTED when the plan does not run the persistent row kernel (small batches): use
 * trt_dft_plan_set_peer_row_split on the column pass instead.
 */
int trt_dft_plan_set_peer_column_gather(trt_dft_plan_t plan, int nparts, const int64_t* col_starts, int64_t rows_per_image,
                                        int64_t image_rows, int64_t row_offset, void* const* peer_slabs, const int64_t* slab_pitches);

/*
 * The mirror image for the pencil irfft2: a trt_dft_plan_create_c2c plan over data[outer][n][inner] stores output row
 * q of every image into the row band of the rank that owns it,
 *     peer_bands[j][image][row_starts[j+1] - row_starts[j]][dest_pitch]   for row_starts[j] <= q < row_starts[j+1],
 * at columns dest_col_offset .. dest_col_offset + inner - 1, again with ordinary stores to peer memory; `out` is then
 * unused.  row_starts has nparts+1 entries from 0 to n.  nparts = 0 restores the ordinary output.
 */
int trt_dft_plan_set_peer_row_split(trt_dft_plan_t plan, int nparts, const int64_t* row_starts, int64_t dest_pitch,
                                    int64_t dest_col_offset, void* const* peer_bands);

/* Replaces the workSize out-parameter of cufftXtMakePlanMany (dft_plugins.cpp:175) that
 * getWorkspaceSize reports (:114-129).  0 for every single-kernel plan. */
size_t trt_dft_workspace_bytes(trt_dft_plan_t plan);

/* Host-only variant of the above for TensorRT's getWorkspaceSize (dft_plugins.cpp:114-129), which is
 * asked about shapes before any plan exists: runs the planner without touching a device.
 * Returns 0 for invalid arguments too. */
size_t trt_dft_query_workspace_bytes(int direction, int rank, const int64_t* dims, int64_t batch);

/*
 * Replaces cufftSetStream + cufftSetWorkArea + cufftXtExec (+ cublasSetStream + cublasScalEx
 * for the inverse): dft_plugins.cpp:185-195 and :454-468.
 *   in, out   : device pointers (out-of-place; `in` is never written -- the reference
 *               const_casts it for cuFFT, :193)
 *   workspace : device pointer to >= trt_dft_workspace_bytes(plan) bytes, 16-byte aligned;
 *               may be NULL when that is 0
 *   stream    : a cudaStream_t (0 = legacy default stream)
 * Asynchronous; no allocation, no synchronisation, no host<->device copy.
 */
int trt_dft_exec(trt_dft_plan_t plan, const void* in, void* out, void* workspace, void* stream);

/* Replaces cufftDestroy / cublasDestroy (dft_plugins.cpp:277-303).  Frees the plan's device
 * tables (synchronises the device, like cudaFree does).  NULL is accepted. */
int trt_dft_plan_destroy(trt_dft_plan_t plan);

/*
 * Host-buffer convenience used by tests and bench.py's end-to-end number: what
 * tests/test_dft.py:104-115 (run_trt_inference) does around the engine -- device buffers,
 * host->device copy of `host_in`, exec, device->host copy into `host_out`, stream
 * synchronise.  `host_in` / `host_out` are ordinary or pinned host memory.  Allocates and
 * frees its own device buffers and workspace.  Synchronous.
 */
int trt_dft_exec_host(trt_dft_plan_t plan, const void* host_in, void* host_out);

/* Introspection for tests and DESIGN.md: fills `buf` (size `len`) with a one-line description
 * of the kernels the plan runs, e.g. "fused2d<90,180> tiles=6144 threads=256 smem=64816".
 * Returns the number of kernel launches one trt_dft_exec issues (> 0), or -1. */
int trt_dft_plan_describe(trt_dft_plan_t plan, char* buf, size_t len);

/*
 * No counterpart in the reference (its cuFFT work area always comes from TensorRT's workspace,
 * dft_plugins.cpp:189).  Same as trt_dft_exec, for callers that cannot supply a workspace
 * buffer because they cannot allocate one where they run -- inside an enqueue, in a serving
 * loop that must not hit cudaMalloc's device-wide synchronisation, or behind an op interface
 * that has inputs and outputs only.  The workspace is a fixed region of
 * TRT_DFT_STATIC_WS_BYTES: a __device__ array, which a GPU holds from the first time this entry
 * point runs on it (that call's address lookup makes the runtime load the array's module there)
 * until the process exits; the call itself allocates and frees nothing.
 *   - plans needing more than TRT_DFT_STATIC_WS_BYTES are refused with
 *     TRT_DFT_ERR_UNSUPPORTED (and a message on stderr); plans needing none run as usual;
 *   - all calls on one GPU share the region: calls whose GPU work may overlap (different
 *     streams) race -- keep them on one stream or order them with events;
 *   - the region is never cleared and never needs to be: each call overwrites what it reads;
 *   - the first call on each GPU looks up the region's device address (with that GPU current);
 *     make that first call before beginning a stream capture on that GPU.
 */
#define TRT_DFT_STATIC_WS_BYTES (32u * 1024u * 1024u)
int trt_dft_exec_static_ws(trt_dft_plan_t plan, const void* in, void* out, void* stream);

/* Number of elements (floats) of the input / output tensors of a plan, for callers that
 * size buffers from the plan instead of recomputing the shape rules. */
int64_t trt_dft_plan_input_floats(trt_dft_plan_t plan);
int64_t trt_dft_plan_output_floats(trt_dft_plan_t plan);

#ifdef __cplusplus
}
#endif
#endif /* TRT_DFT_H_ */
