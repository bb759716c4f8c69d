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
//   * plans are looked up from the descriptors passed to each call, so a clone() or a
//     deserialised plugin works even before configurePlugin ran on that instance (the reference's
//     clones are unusable until then, :73-83), and getWorkspaceSize is derived from its arguments.
//
// Compiled against the real <NvInfer.h> when TENSORRT_INCLUDE_DIR is given to build.py; otherwise
// against the stand-in trt_min/NvInfer.h (test aid -- see the warning in that file).
#include <NvInfer.h>

#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <map>
#include <mutex>
#include <new>
#include <string>
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

class DftPlugin final : public IPluginV2DynamicExt {
public:
    DftPlugin(Op op, Attributes a) : op_(op), attrs_(a) {}
    ~DftPlugin() noexcept override
    {
        for (auto& kv : plans_) trt_dft_plan_destroy(kv.second);
    }

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
        return inOut[pos].format == TensorFormat::kLINEAR && inOut[pos].type == DataType::kFLOAT;
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
        if (p) p->ns_ = ns_;
        return p;
    }

    void configurePlugin(DynamicPluginTensorDesc const* in, int32_t nbInputs, DynamicPluginTensorDesc const* out,
                         int32_t nbOutputs) noexcept override
    {
        // Build the plan for the configured shape now, so that enqueue only launches.
        if (nbInputs != 1 || nbOutputs != 1 || !in || !out) return;
        if (in[0].desc.type != DataType::kFLOAT || out[0].desc.type != DataType::kFLOAT) return;
        (void)plan_for(in[0].desc, out[0].desc);
    }

    size_t getWorkspaceSize(PluginTensorDesc const* inputs, int32_t nbInputs, PluginTensorDesc const* outputs,
                            int32_t nbOutputs) const noexcept override
    {
        if (nbInputs != 1 || nbOutputs != 1 || !inputs || !outputs) return 0;
        int64_t batch = 0, dims[3] = {1, 1, 1};
        if (!split_dims(signal_desc(inputs[0], outputs[0]).dims, attrs_.signal_ndim, &batch, dims)) return 0;
        return trt_dft_query_workspace_bytes(direction(), attrs_.signal_ndim, dims, batch);   // host-only arithmetic
    }

    int32_t enqueue(PluginTensorDesc const* inputDesc, PluginTensorDesc const* outputDesc, void const* const* inputs,
                    void* const* outputs, void* workspace, cudaStream_t stream) noexcept override
    {
        if (!inputDesc || !outputDesc || !inputs || !outputs) return 1;
        trt_dft_plan_t plan = plan_for(inputDesc[0], outputDesc[0]);
        if (!plan) return 1;
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

    trt_dft_plan_t plan_for(PluginTensorDesc const& in, PluginTensorDesc const& out) noexcept
    {
        int64_t batch = 0, dims[3] = {1, 1, 1};
        if (!split_dims(signal_desc(in, out).dims, attrs_.signal_ndim, &batch, dims)) return nullptr;
        std::vector<int64_t> key{batch, dims[0], dims[1], dims[2]};
        std::lock_guard<std::mutex> lock(mu_);
        try {
            auto it = plans_.find(key);
            if (it != plans_.end()) return it->second;
            trt_dft_plan_t p = nullptr;
            if (trt_dft_plan_create(&p, direction(), attrs_.signal_ndim, dims, batch) != TRT_DFT_OK) return nullptr;
            plans_[key] = p;
            return p;
        } catch (...) {
            return nullptr;
        }
    }

    Op op_;
    Attributes attrs_;
    std::string ns_;
    std::mutex mu_;
    std::map<std::vector<int64_t>, trt_dft_plan_t> plans_;
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
// Load-time registration, the behaviour the reference gets from its two static PluginRegistrar
// objects (:575-576): after dlopen, "Rfft" and "Irfft" are in the global registry under namespace "".
struct RegisterAtLoad {
    RegisterAtLoad()
    {
        if (!getPluginRegistry) return;
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
    nbCreators = 2;
    return creators;
}
extern "C" __attribute__((visibility("default"))) void setLoggerFinder(void*) {}
