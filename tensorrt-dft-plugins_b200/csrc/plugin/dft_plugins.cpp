// dft_plugins.cpp -- the "Rfft" / "Irfft" TensorRT plugin shell over the C ABI of trt_dft.h.
//
// Observable behaviour follows the reference's plugin boundary
// (/root/reference/src/dft_plugins/dft_plugins.cpp): plugin names "Rfft" / "Irfft" (:385, :475),
// version "1" (:43), namespace "", three INT32 creator fields normalized / onesided / signal_ndim in
// that order (:500-507), accepted values normalized == 0, onesided == 1, 1 <= signal_ndim <= 3
// (:55-57), one fp32 linear input and one fp32 linear output (:93-105, :232-240), the shape rules
// of :361-382 and :415-436, and the 12-byte serialisation of :201-218.  What differs on purpose:
//   * enqueue is a few lines over trt_dft_exec -- no cuFFT, no cuBLAS, no separate scale pass;
//   * nothing asserts: bad attributes / blobs make the creator return nullptr, failures in
//     enqueue return non-zero;
//   * plans live in one process-wide cache keyed by (op, signal_ndim, dims, batch, device), filled by
//     configurePlugin (where the reference builds its cuFFT plan, :131-178).  A clone() or a
//     deserialised plugin therefore finds the plan of the instance that was configured, and its first
//     enqueue is a pure lookup (the reference's clones are unusable until configured, :73-83).  enqueue
//     itself never allocates while the stream is being captured: a shape nobody configured is then an
//     error instead of a cudaMalloc inside the capture;
//   * getWorkspaceSize is derived from its arguments.
//
// Compiled against the real <NvInfer.h> when TENSORRT_INCLUDE_DIR is given to build.py; otherwise
// against the stand-in trt_min/NvInfer.h (test aid -- see the warning in that file).
#include <NvInfer.h>

#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <cuda_runtime_api.h>
#include <dlfcn.h>

#include <map>
#include <mutex>
#include <new>
#include <string>
#include <tuple>
#include <vector>

#include "../../../include/trt_dft.h"

using namespace nvinfer1;

namespace trt_dft {

namespace {

constexpr char kPluginVersion[] = "1";
constexpr char const* kFieldNames[3] = {"normalized", "onesided", "signal_ndim"};

struct Attributes {
    int32_t normalized;
    int32_t onesided;
    int32_t signal_ndim;
    bool acceptable() const { return normalized == 0 && onesided == 1 && signal_ndim >= 1 && signal_ndim <= 3; }
};
static_assert(sizeof(Attributes) == 12, "wire format is three host-endian int32");

enum class Op { kRfft, kIrfft };
char const* op_name(Op op) { return op == Op::kRfft ? "Rfft" : "Irfft"; }

// batch and signal dims from a tensor's dims: the trailing signal_ndim dims are transformed, all
// leading dims fold into the batch (reference: splitSignalDims, :250-266).
bool split_dims(Dims const& d, int32_t signal_ndim, int64_t* batch, int64_t out[3])
{
    if (d.nbDims < signal_ndim) return false;
    int64_t b = 1;
    const int32_t first = d.nbDims - signal_ndim;
    for (int32_t i = 0; i < first; ++i) b *= d.d[i];
    for (int32_t i = 0; i < signal_ndim; ++i) out[i] = d.d[first + i];
    *batch = b;
    return true;
}

// Fixed-size key of one plan: no heap traffic on the enqueue path.
struct PlanKey {
    int32_t direction, signal_ndim, device, io_type;
    int64_t batch, d0, d1, d2;
    auto tie() const { return std::tie(direction, signal_ndim, device, io_type, batch, d0, d1, d2); }
    bool operator<(PlanKey const& o) const { return tie() < o.tie(); }
    bool operator==(PlanKey const& o) const { return tie() == o.tie(); }
};

// Process-wide plan cache.  Plans are immutable once built and usable from any stream, so every plugin
// instance (configured original, builder clone, deserialised runtime copy, per-context clone) shares them.
// Entries stay until the process exits: a plan is a few KB of tables, and a handle that enqueue may be using
// on another thread must not disappear under it.
class PlanCache {
public:
    static PlanCache& get()
    {
        static PlanCache* c = new PlanCache();     // never destroyed: outlives the CUDA runtime's own teardown
        return *c;
    }
    // may_create = false: lookup only (enqueue during stream capture)
    trt_dft_plan_t find_or_create(PlanKey const& k, bool may_create) noexcept
    {
        std::lock_guard<std::mutex> lock(mu_);
        try {
            auto it = plans_.find(k);
            if (it != plans_.end()) return it->second;
            if (!may_create) return nullptr;
            int64_t dims[3] = {k.d0, k.d1, k.d2};
            trt_dft_plan_t p = nullptr;
            if (trt_dft_plan_create_typed(&p, k.direction, k.signal_ndim, dims, k.batch, k.io_type) != TRT_DFT_OK) return nullptr;
            plans_[k] = p;
            return p;
        } catch (...) {
            return nullptr;
        }
    }

private:
    std::mutex mu_;
    std::map<PlanKey, trt_dft_plan_t> plans_;
};

// TensorRT data type -> TRT_DFT_IO_*, or -1.  The reference takes fp32 only (dft_plugins.cpp:93-105) and this plugin does
// the same unless TRT_DFT_PLUGIN_16BIT_IO=1 is set when the engine is BUILT: with it, fp16 (and, from TensorRT 9, bf16)
// tensors are offered as well -- same type on both sides, fp32 arithmetic inside -- which saves a half-precision engine
// the two conversion layers around the plugin.  Opt-in because it changes numerics a model may rely on: the outputs are
// rounded to 16 bits, and an unnormalised forward transform of a large image can exceed fp16's range (65504).
static int io_type_of(DataType t)
{
    if (t == DataType::kFLOAT) return TRT_DFT_IO_F32;
    char const* e = getenv("TRT_DFT_PLUGIN_16BIT_IO");        // (only ever read for non-fp32 tensors)
    if (!e || e[0] != '1') return -1;
    if (t == DataType::kHALF) return TRT_DFT_IO_F16;
#if NV_TENSORRT_MAJOR >= 9
    if (t == DataType::kBF16) return TRT_DFT_IO_BF16;
#endif
    return -1;
}

class DftPlugin final : public IPluginV2DynamicExt {
public:
    DftPlugin(Op op, Attributes a) : op_(op), attrs_(a) {}
    ~DftPlugin() noexcept override = default;

    // --- identity ---------------------------------------------------------------------------------
    AsciiChar const* getPluginType() const noexcept override { return op_name(op_); }
    AsciiChar const* getPluginVersion() const noexcept override { return kPluginVersion; }
    int32_t getNbOutputs() const noexcept override { return 1; }
    void setPluginNamespace(AsciiChar const* ns) noexcept override { ns_ = ns ? ns : ""; }
    AsciiChar const* getPluginNamespace() const noexcept override { return ns_.c_str(); }

    // --- shape / type / format negotiation ----------------------------------------------------------
    DimsExprs getOutputDimensions(int32_t outputIndex, DimsExprs const* inputs, int32_t nbInputs,
                                  IExprBuilder& eb) noexcept override
    {
        DimsExprs out{};
        if (outputIndex != 0 || nbInputs != 1 || !inputs) return out;
        out = inputs[0];
        if (op_ == Op::kRfft) {
            // [..., n] -> [..., n/2 + 1, 2]
            if (out.nbDims < 1 || out.nbDims >= Dims::MAX_DIMS) return DimsExprs{};
            IDimensionExpr const* last = out.d[out.nbDims - 1];
            out.d[out.nbDims - 1] =
                eb.operation(DimensionOperation::kSUM, *eb.operation(DimensionOperation::kFLOOR_DIV, *last, *eb.constant(2)),
                             *eb.constant(1));
            out.d[out.nbDims] = eb.constant(2);
            out.nbDims += 1;
        } else {
            // [..., m, 2] -> [..., (m - 1) * 2]
            if (out.nbDims < 2) return DimsExprs{};
            out.nbDims -= 1;
            IDimensionExpr const* m = out.d[out.nbDims - 1];
            out.d[out.nbDims - 1] =
                eb.operation(DimensionOperation::kPROD, *eb.operation(DimensionOperation::kSUB, *m, *eb.constant(1)),
                             *eb.constant(2));
        }
        return out;
    }

    bool supportsFormatCombination(int32_t pos, PluginTensorDesc const* inOut, int32_t nbInputs,
                                   int32_t nbOutputs) noexcept override
    {
        if (nbInputs != 1 || nbOutputs != 1 || pos < 0 || pos >= 2 || !inOut) return false;
        if (inOut[pos].format != TensorFormat::kLINEAR || io_type_of(inOut[pos].type) < 0) return false;
        return pos == 0 || inOut[1].type == inOut[0].type;          // one type on both sides
    }

    DataType getOutputDataType(int32_t, DataType const* inputTypes, int32_t nbInputs) const noexcept override
    {
        return (nbInputs == 1 && inputTypes) ? inputTypes[0] : DataType::kFLOAT;
    }

    // --- lifecycle ------------------------------------------------------------------------------------
    int32_t initialize() noexcept override { return 0; }
    void terminate() noexcept override {}
    void destroy() noexcept override { delete this; }

    IPluginV2DynamicExt* clone() const noexcept override
    {
        DftPlugin* p = new (std::nothrow) DftPlugin(op_, attrs_);
        if (p) {
            p->ns_ = ns_;
            p->dynamic_ = dynamic_;
            p->io_type_ = io_type_;
        }
        return p;
    }

    void configurePlugin(DynamicPluginTensorDesc const* in, int32_t nbInputs, DynamicPluginTensorDesc const* out,
                         int32_t nbOutputs) noexcept override
    {
        // Build the plan for the configured shape now, so that enqueue only launches.
        if (nbInputs != 1 || nbOutputs != 1 || !in || !out) return;
        if (io_type_of(in[0].desc.type) < 0 || out[0].desc.type != in[0].desc.type) return;
        io_type_ = io_type_of(in[0].desc.type);
        // With a dynamic-shape profile (min != max) build the plans of the profile's corner shapes as well, so
        // that those shapes never allocate in enqueue; shapes in between are planned on first use.
        dynamic_ = false;
        Dims const& smin = op_ == Op::kRfft ? in[0].min : out[0].min;
        Dims const& smax = op_ == Op::kRfft ? in[0].max : out[0].max;
        if (smin.nbDims == smax.nbDims)
            for (int32_t i = 0; i < smin.nbDims; ++i) dynamic_ = dynamic_ || smin.d[i] != smax.d[i];
        (void)plan_for(signal_desc(in[0].desc, out[0].desc).dims, /*may_create=*/true);
        if (dynamic_) {
            (void)plan_for(smin, true);
            (void)plan_for(smax, true);
        }
    }

    size_t getWorkspaceSize(PluginTensorDesc const* inputs, int32_t nbInputs, PluginTensorDesc const* outputs,
                            int32_t nbOutputs) const noexcept override
    {
        if (nbInputs != 1 || nbOutputs != 1 || !inputs || !outputs) return 0;
        int64_t batch = 0, dims[3] = {1, 1, 1};
        if (!split_dims(signal_desc(inputs[0], outputs[0]).dims, attrs_.signal_ndim, &batch, dims)) return 0;
        size_t need = trt_dft_query_workspace_bytes(direction(), attrs_.signal_ndim, dims, batch);   // host-only arithmetic
        const bool io16 = io_type_of(inputs[0].type) > TRT_DFT_IO_F32;
        if (dynamic_) {
            // TensorRT asks for the profile's largest shape only, but a smaller shape may pick another kind of
            // plan (fused kernels need nothing, un-fused inverses one complex tensor, lengths outside the radix
            // set two).  Two complex tensors of the largest shape bound every plan of every smaller shape.
            int64_t cplx = batch * 8;
            for (int32_t i = 0; i < attrs_.signal_ndim; ++i) cplx *= (i == attrs_.signal_ndim - 1) ? dims[i] / 2 + 1 : dims[i];
            const size_t bound = 2 * (((size_t)cplx + 255) / 256 * 256);
            if (bound > need) need = bound;
        }
        if (io16) {
            // shapes without converting kernels run the fp32 plan between two conversion passes: fp32 copies of both
            // tensors ahead of the plan's own workspace (an upper bound for the shapes that need none of it)
            int64_t real = batch, cplx = batch * 2;
            for (int32_t i = 0; i < attrs_.signal_ndim; ++i) {
                real *= dims[i];
                cplx *= (i == attrs_.signal_ndim - 1) ? dims[i] / 2 + 1 : dims[i];
            }
            need += ((size_t)real * 4 + 255) / 256 * 256 + ((size_t)cplx * 4 + 255) / 256 * 256;
        }
        return need;
    }

    int32_t enqueue(PluginTensorDesc const* inputDesc, PluginTensorDesc const* outputDesc, void const* const* inputs,
                    void* const* outputs, void* workspace, cudaStream_t stream) noexcept override
    {
        if (!inputDesc || !outputDesc || !inputs || !outputs) return 1;
        Dims const& sd = signal_desc(inputDesc[0], outputDesc[0]).dims;
        const int io = io_type_of(inputDesc[0].type);
        if (io < 0 || outputDesc[0].type != inputDesc[0].type) return 1;
        io_type_ = io;
        trt_dft_plan_t plan = plan_for(sd, /*may_create=*/false);       // the normal case: a lookup, nothing else
        if (!plan) {
            // A shape no instance was configured for (dynamic shapes between the profile's corners).  Building a
            // plan allocates and copies tables, which is not allowed inside a stream capture: refuse there.
            cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
            if (cudaStreamIsCapturing(stream, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) {
                cudaGetLastError();
                fprintf(stderr, "[trt_dft_plugins] %s enqueue: no plan for this shape and the stream is being captured; "
                                "run the shape once outside the capture (or configure it) first\n", op_name(op_));
                return 1;
            }
            plan = plan_for(sd, /*may_create=*/true);
            if (!plan) return 1;
        }
        return trt_dft_exec(plan, inputs[0], outputs[0], workspace, stream) == TRT_DFT_OK ? 0 : 1;
    }

    // --- serialisation: normalized, onesided, signal_ndim as three int32 ---------------------------------
    size_t getSerializationSize() const noexcept override { return sizeof(Attributes); }
    void serialize(void* buffer) const noexcept override
    {
        if (buffer) memcpy(buffer, &attrs_, sizeof(attrs_));
    }

private:
    int direction() const { return op_ == Op::kRfft ? TRT_DFT_FORWARD : TRT_DFT_INVERSE; }
    // the tensor that carries the logical real size: the input of Rfft, the output of Irfft
    PluginTensorDesc const& signal_desc(PluginTensorDesc const& in, PluginTensorDesc const& out) const
    {
        return op_ == Op::kRfft ? in : out;
    }

    // signal dims -> plan.  The instance remembers the last plan it used (TensorRT drives one instance from one
    // thread at a time), so a steady-state enqueue touches neither the shared cache nor its mutex.
    trt_dft_plan_t plan_for(Dims const& signal_dims, bool may_create) noexcept
    {
        int64_t batch = 0, dims[3] = {1, 1, 1};
        if (!split_dims(signal_dims, attrs_.signal_ndim, &batch, dims)) return nullptr;
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) {
            cudaGetLastError();
            return nullptr;
        }
        const PlanKey key{direction(), attrs_.signal_ndim, dev, io_type_, batch, dims[0], dims[1], dims[2]};
        if (last_plan_ && key == last_key_) return last_plan_;
        trt_dft_plan_t p = PlanCache::get().find_or_create(key, may_create);
        if (p) {
            last_key_ = key;
            last_plan_ = p;
        }
        return p;
    }

    Op op_;
    Attributes attrs_;
    std::string ns_;
    bool dynamic_ = false;              // configured with a dynamic-shape profile (min != max)
    int32_t io_type_ = TRT_DFT_IO_F32;  // of the tensors last configured / enqueued
    PlanKey last_key_{};
    trt_dft_plan_t last_plan_ = nullptr;
};

class DftPluginCreator final : public IPluginCreator {
public:
    explicit DftPluginCreator(Op op) : op_(op)
    {
        for (char const* n : kFieldNames) fields_.emplace_back(PluginField{n, nullptr, PluginFieldType::kINT32, 1});
        collection_.nbFields = (int32_t)fields_.size();
        collection_.fields = fields_.data();
    }
    AsciiChar const* getPluginName() const noexcept override { return op_name(op_); }
    AsciiChar const* getPluginVersion() const noexcept override { return kPluginVersion; }
    PluginFieldCollection const* getFieldNames() noexcept override { return &collection_; }

    IPluginV2* createPlugin(AsciiChar const*, PluginFieldCollection const* fc) noexcept override
    {
        if (!fc || !fc->fields) return complain("no field collection");
        int32_t v[3] = {0, 0, 0};
        bool seen[3] = {false, false, false};
        for (int32_t i = 0; i < fc->nbFields; ++i) {
            PluginField const& f = fc->fields[i];
            if (!f.name) continue;
            for (int j = 0; j < 3; ++j) {
                if (strcmp(f.name, kFieldNames[j]) != 0) continue;
                if (f.type != PluginFieldType::kINT32 || !f.data || f.length < 1) return complain("fields must be INT32 scalars");
                memcpy(&v[j], f.data, sizeof(int32_t));
                seen[j] = true;
            }
        }
        if (!(seen[0] && seen[1] && seen[2])) return complain("needs normalized, onesided and signal_ndim");
        return make(Attributes{v[0], v[1], v[2]});
    }

    IPluginV2* deserializePlugin(AsciiChar const*, void const* data, size_t length) noexcept override
    {
        if (!data || length != sizeof(Attributes)) return complain("serialised blob must be 12 bytes");
        Attributes a;
        memcpy(&a, data, sizeof(a));
        return make(a);
    }

    void setPluginNamespace(AsciiChar const* ns) noexcept override { ns_ = ns ? ns : ""; }
    AsciiChar const* getPluginNamespace() const noexcept override { return ns_.c_str(); }

private:
    IPluginV2* make(Attributes a) noexcept
    {
        // the com.microsoft Rfft / Irfft contrib ops only define this combination (reference :54-57)
        if (!a.acceptable()) return complain("unsupported attributes (need normalized=0, onesided=1, signal_ndim in 1..3)");
        DftPlugin* p = new (std::nothrow) DftPlugin(op_, a);
        if (p) p->setPluginNamespace(ns_.c_str());
        return p;
    }
    IPluginV2* complain(char const* what) noexcept
    {
        fprintf(stderr, "[trt_dft_plugins] %s plugin: %s\n", op_name(op_), what);
        return nullptr;
    }

    Op op_;
    std::vector<PluginField> fields_;
    PluginFieldCollection collection_{};
    std::string ns_;
};

DftPluginCreator& rfft_creator()
{
    static DftPluginCreator c(Op::kRfft);
    return c;
}
DftPluginCreator& irfft_creator()
{
    static DftPluginCreator c(Op::kIrfft);
    return c;
}

}  // namespace
}  // namespace trt_dft

// libnvinfer's registry accessor, resolved lazily: the library must also load in a process that
// has no TensorRT (the C ABI is usable on its own).
extern "C" nvinfer1::IPluginRegistry* getPluginRegistry() noexcept __attribute__((weak));

namespace {
// A library compiled against the stand-in header (no TensorRT at build time) has plugin classes whose
// virtual-table layout was never checked against a real libnvinfer.  Handing such objects to a real
// TensorRT would crash or misdispatch, so in that build the creators are only given to a host that
// identifies itself as the test double (tests/mock_trt exports `trt_dft_mock_host`).
bool host_may_have_creators()
{
#if defined(TRT_MIN_STAND_IN)
    if (dlsym(RTLD_DEFAULT, "trt_dft_mock_host") != nullptr) return true;
    static bool warned = false;
    if (!warned && getPluginRegistry) {
        warned = true;
        fprintf(stderr,
                "[trt_dft_plugins] this libtrt_dft_plugins.so was built WITHOUT TensorRT headers (stand-in NvInfer.h); "
                "the Rfft / Irfft creators are not registered.  Rebuild with TENSORRT_INCLUDE_DIR=<dir with NvInfer.h>.\n");
    }
    return false;
#else
    return true;
#endif
}

// Load-time registration, the behaviour the reference gets from its two static PluginRegistrar
// objects (:575-576): after dlopen, "Rfft" and "Irfft" are in the global registry under namespace "".
struct RegisterAtLoad {
    RegisterAtLoad()
    {
        if (!getPluginRegistry) return;
        if (!host_may_have_creators()) return;
        nvinfer1::IPluginRegistry* reg = getPluginRegistry();
        if (!reg) return;
        reg->registerCreator(trt_dft::rfft_creator(), "");
        reg->registerCreator(trt_dft::irfft_creator(), "");
    }
} g_register_at_load;
}  // namespace

// For hosts that load plugin libraries through IPluginRegistry::loadLibrary (TensorRT >= 8.6), which
// asks the library for its creators instead of relying on static initialisers.
extern "C" __attribute__((visibility("default"))) nvinfer1::IPluginCreator* const* getPluginCreators(int32_t& nbCreators)
{
    static nvinfer1::IPluginCreator* const creators[2] = {&trt_dft::rfft_creator(), &trt_dft::irfft_creator()};
    nbCreators = host_may_have_creators() ? 2 : 0;
    return creators;
}
extern "C" __attribute__((visibility("default"))) void setLoggerFinder(void*) {}
