// trt_min/NvInfer.h -- MINIMAL STAND-IN for TensorRT's public plugin interfaces.
//
// TensorRT is not installed in this image (no NvInfer.h, no libnvinfer).  This header declares,
// from the public TensorRT 8.x API documentation as recalled, only the types the Rfft / Irfft
// plugin shell implements, so that the shell can be compiled and driven by the mock host in
// tests/mock_trt.  It is a test aid, NOT an ABI guarantee: virtual-method order, struct layouts and
// enum values must be re-verified against a real NvInfer.h, and build.py prefers the real header
// whenever one is found (set TENSORRT_INCLUDE_DIR).  Nothing here is copied from TensorRT sources.
#ifndef TRT_MIN_NVINFER_H_
#define TRT_MIN_NVINFER_H_

#include <cuda_runtime_api.h>
#include <stddef.h>
#include <stdint.h>

#define TRT_MIN_STAND_IN 1
#define NV_TENSORRT_MAJOR 8
#define NV_TENSORRT_MINOR 6
#define NV_TENSORRT_PATCH 1
#define NV_TENSORRT_VERSION (NV_TENSORRT_MAJOR * 1000 + NV_TENSORRT_MINOR * 100 + NV_TENSORRT_PATCH)

namespace nvinfer1 {

using AsciiChar = char;

enum class DataType : int32_t { kFLOAT = 0, kHALF = 1, kINT8 = 2, kINT32 = 3, kBOOL = 4, kUINT8 = 5, kFP8 = 6 };
enum class TensorFormat : int32_t { kLINEAR = 0, kCHW2 = 1, kHWC8 = 2, kCHW4 = 3, kCHW16 = 4, kCHW32 = 5 };
enum class PluginFieldType : int32_t {
    kFLOAT16 = 0, kFLOAT32 = 1, kFLOAT64 = 2, kINT8 = 3, kINT16 = 4, kINT32 = 5, kCHAR = 6, kDIMS = 7, kUNKNOWN = 8
};
enum class DimensionOperation : int32_t {
    kSUM = 0, kPROD = 1, kMAX = 2, kMIN = 3, kSUB = 4, kEQUAL = 5, kLESS = 6, kFLOOR_DIV = 7, kCEIL_DIV = 8
};

class Dims {
public:
    static constexpr int32_t MAX_DIMS{8};
    int32_t nbDims;
    int32_t d[MAX_DIMS];
};

class IDimensionExpr {
public:
    virtual bool isConstant() const noexcept = 0;
    virtual int32_t getConstantValue() const noexcept = 0;
protected:
    virtual ~IDimensionExpr() noexcept = default;
};

class DimsExprs {
public:
    int32_t nbDims;
    IDimensionExpr const* d[Dims::MAX_DIMS];
};

class IExprBuilder {
public:
    virtual IDimensionExpr const* constant(int32_t value) noexcept = 0;
    virtual IDimensionExpr const* operation(DimensionOperation op, IDimensionExpr const& first,
                                            IDimensionExpr const& second) noexcept = 0;
protected:
    virtual ~IExprBuilder() noexcept = default;
};

struct PluginTensorDesc {
    Dims dims;
    DataType type;
    TensorFormat format;
    float scale;
};

struct DynamicPluginTensorDesc {
    PluginTensorDesc desc;
    Dims min;
    Dims max;
};

class PluginField {
public:
    AsciiChar const* name;
    void const* data;
    PluginFieldType type;
    int32_t length;
    PluginField(AsciiChar const* const name_ = nullptr, void const* const data_ = nullptr,
                PluginFieldType const type_ = PluginFieldType::kUNKNOWN, int32_t const length_ = 0) noexcept
        : name(name_), data(data_), type(type_), length(length_) {}
};

struct PluginFieldCollection {
    int32_t nbFields;
    PluginField const* fields;
};

enum class PluginFormat : int32_t { kLINEAR = 0 };
class IGpuAllocator;

class IPluginV2 {
public:
    virtual int32_t getTensorRTVersion() const noexcept { return NV_TENSORRT_VERSION; }
    virtual AsciiChar const* getPluginType() const noexcept = 0;
    virtual AsciiChar const* getPluginVersion() const noexcept = 0;
    virtual int32_t getNbOutputs() const noexcept = 0;
    virtual Dims getOutputDimensions(int32_t index, Dims const* inputs, int32_t nbInputDims) noexcept = 0;
    virtual bool supportsFormat(DataType type, PluginFormat format) const noexcept = 0;
    virtual void configureWithFormat(Dims const* inputDims, int32_t nbInputs, Dims const* outputDims, int32_t nbOutputs,
                                     DataType type, PluginFormat format, int32_t maxBatchSize) noexcept = 0;
    virtual int32_t initialize() noexcept = 0;
    virtual void terminate() noexcept = 0;
    virtual size_t getWorkspaceSize(int32_t maxBatchSize) const noexcept = 0;
    virtual int32_t enqueue(int32_t batchSize, void const* const* inputs, void* const* outputs, void* workspace,
                            cudaStream_t stream) noexcept = 0;
    virtual size_t getSerializationSize() const noexcept = 0;
    virtual void serialize(void* buffer) const noexcept = 0;
    virtual void destroy() noexcept = 0;
    virtual IPluginV2* clone() const noexcept = 0;
    virtual void setPluginNamespace(AsciiChar const* pluginNamespace) noexcept = 0;
    virtual AsciiChar const* getPluginNamespace() const noexcept = 0;
protected:
    virtual ~IPluginV2() noexcept = default;
};

class IPluginV2Ext : public IPluginV2 {
public:
    virtual DataType getOutputDataType(int32_t index, DataType const* inputTypes, int32_t nbInputs) const noexcept = 0;
    virtual bool isOutputBroadcastAcrossBatch(int32_t outputIndex, bool const* inputIsBroadcasted,
                                              int32_t nbInputs) const noexcept = 0;
    virtual bool canBroadcastInputAcrossBatch(int32_t inputIndex) const noexcept = 0;
    virtual void configurePlugin(Dims const* inputDims, int32_t nbInputs, Dims const* outputDims, int32_t nbOutputs,
                                 DataType const* inputTypes, DataType const* outputTypes, bool const* inputIsBroadcast,
                                 bool const* outputIsBroadcast, PluginFormat floatFormat, int32_t maxBatchSize) noexcept = 0;
    virtual void attachToContext(void* /*cudnn*/, void* /*cublas*/, IGpuAllocator* /*allocator*/) noexcept {}
    virtual void detachFromContext() noexcept {}
    IPluginV2Ext* clone() const noexcept override = 0;
};

class IPluginV2DynamicExt : public IPluginV2Ext {
public:
    IPluginV2DynamicExt* clone() const noexcept override = 0;
    virtual DimsExprs getOutputDimensions(int32_t outputIndex, DimsExprs const* inputs, int32_t nbInputs,
                                          IExprBuilder& exprBuilder) noexcept = 0;
    virtual bool supportsFormatCombination(int32_t pos, PluginTensorDesc const* inOut, int32_t nbInputs,
                                           int32_t nbOutputs) noexcept = 0;
    virtual void configurePlugin(DynamicPluginTensorDesc const* in, int32_t nbInputs,
                                 DynamicPluginTensorDesc const* out, int32_t nbOutputs) noexcept = 0;
    virtual size_t getWorkspaceSize(PluginTensorDesc const* inputs, int32_t nbInputs, PluginTensorDesc const* outputs,
                                    int32_t nbOutputs) const noexcept = 0;
    virtual int32_t enqueue(PluginTensorDesc const* inputDesc, PluginTensorDesc const* outputDesc,
                            void const* const* inputs, void* const* outputs, void* workspace,
                            cudaStream_t stream) noexcept = 0;
protected:
    // the batch-era overloads of IPluginV2 / IPluginV2Ext are not used by dynamic-shape plugins
    Dims getOutputDimensions(int32_t, Dims const*, int32_t) noexcept override final { return Dims{-1, {}}; }
    bool isOutputBroadcastAcrossBatch(int32_t, bool const*, int32_t) const noexcept override final { return false; }
    bool canBroadcastInputAcrossBatch(int32_t) const noexcept override final { return true; }
    bool supportsFormat(DataType, PluginFormat) const noexcept override final { return false; }
    void configureWithFormat(Dims const*, int32_t, Dims const*, int32_t, DataType, PluginFormat,
                             int32_t) noexcept override final {}
    void configurePlugin(Dims const*, int32_t, Dims const*, int32_t, DataType const*, DataType const*, bool const*,
                         bool const*, PluginFormat, int32_t) noexcept override final {}
    size_t getWorkspaceSize(int32_t) const noexcept override final { return 0; }
    int32_t enqueue(int32_t, void const* const*, void* const*, void*, cudaStream_t) noexcept override final { return 1; }
};

class IPluginCreator {
public:
    virtual int32_t getTensorRTVersion() const noexcept { return NV_TENSORRT_VERSION; }
    virtual AsciiChar const* getPluginName() const noexcept = 0;
    virtual AsciiChar const* getPluginVersion() const noexcept = 0;
    virtual PluginFieldCollection const* getFieldNames() noexcept = 0;
    virtual IPluginV2* createPlugin(AsciiChar const* name, PluginFieldCollection const* fc) noexcept = 0;
    virtual IPluginV2* deserializePlugin(AsciiChar const* name, void const* serialData, size_t serialLength) noexcept = 0;
    virtual void setPluginNamespace(AsciiChar const* pluginNamespace) noexcept = 0;
    virtual AsciiChar const* getPluginNamespace() const noexcept = 0;
    virtual ~IPluginCreator() = default;
};

class IPluginRegistry {
public:
    virtual bool registerCreator(IPluginCreator& creator, AsciiChar const* const pluginNamespace) noexcept = 0;
    virtual IPluginCreator* const* getPluginCreatorList(int32_t* const numCreators) const noexcept = 0;
    virtual IPluginCreator* getPluginCreator(AsciiChar const* const pluginName, AsciiChar const* const pluginVersion,
                                             AsciiChar const* const pluginNamespace = "") noexcept = 0;
protected:
    virtual ~IPluginRegistry() noexcept = default;
};

}  // namespace nvinfer1

extern "C" nvinfer1::IPluginRegistry* getPluginRegistry() noexcept;

#endif  // TRT_MIN_NVINFER_H_
