// tests/mock_trt/mock_nvinfer.cpp -- TEST DOUBLE for the parts of libnvinfer the plugin library uses.
//
// Provides getPluginRegistry() (so that loading libtrt_dft_plugins.so after this library registers
// the creators exactly as it would under TensorRT), an IExprBuilder that evaluates constant
// expressions, and extern "C" drivers that walk the IPluginV2DynamicExt lifecycle the way the
// TensorRT builder / runtime do: lookup -> createPlugin -> getOutputDimensions ->
// supportsFormatCombination -> clone -> configurePlugin -> getWorkspaceSize -> serialize ->
// deserializePlugin -> clone -> initialize -> enqueue -> terminate -> destroy.
// Test infrastructure only.
#include <NvInfer.h>
#include <cuda_runtime_api.h>
#include <stdio.h>
#include <string.h>

#include <memory>
#include <string>
#include <vector>

using namespace nvinfer1;

namespace {

class Registry final : public IPluginRegistry {
public:
    bool registerCreator(IPluginCreator& creator, AsciiChar const* const ns) noexcept override
    {
        for (auto* c : creators_)
            if (!strcmp(c->getPluginName(), creator.getPluginName()) && !strcmp(c->getPluginVersion(), creator.getPluginVersion()))
                return false;
        creator.setPluginNamespace(ns ? ns : "");
        creators_.push_back(&creator);
        return true;
    }
    IPluginCreator* const* getPluginCreatorList(int32_t* const n) const noexcept override
    {
        *n = (int32_t)creators_.size();
        return creators_.data();
    }
    IPluginCreator* getPluginCreator(AsciiChar const* const name, AsciiChar const* const version,
                                     AsciiChar const* const ns) noexcept override
    {
        for (auto* c : creators_)
            if (!strcmp(c->getPluginName(), name) && !strcmp(c->getPluginVersion(), version) &&
                !strcmp(c->getPluginNamespace(), ns ? ns : ""))
                return c;
        return nullptr;
    }
private:
    std::vector<IPluginCreator*> creators_;
};

Registry g_registry;

struct ConstExpr final : IDimensionExpr {
    int32_t v;
    explicit ConstExpr(int32_t v_) : v(v_) {}
    bool isConstant() const noexcept override { return true; }
    int32_t getConstantValue() const noexcept override { return v; }
};

class Evaluator final : public IExprBuilder {
public:
    IDimensionExpr const* constant(int32_t value) noexcept override
    {
        pool_.emplace_back(new ConstExpr(value));
        return pool_.back().get();
    }
    IDimensionExpr const* operation(DimensionOperation op, IDimensionExpr const& a, IDimensionExpr const& b) noexcept override
    {
        const int32_t x = a.getConstantValue(), y = b.getConstantValue();
        int32_t r = 0;
        switch (op) {
            case DimensionOperation::kSUM: r = x + y; break;
            case DimensionOperation::kPROD: r = x * y; break;
            case DimensionOperation::kMAX: r = x > y ? x : y; break;
            case DimensionOperation::kMIN: r = x < y ? x : y; break;
            case DimensionOperation::kSUB: r = x - y; break;
            case DimensionOperation::kEQUAL: r = x == y; break;
            case DimensionOperation::kLESS: r = x < y; break;
            case DimensionOperation::kFLOOR_DIV: r = x / y; break;
            case DimensionOperation::kCEIL_DIV: r = (x + y - 1) / y; break;
        }
        return constant(r);
    }
private:
    std::vector<std::unique_ptr<ConstExpr>> pool_;
};

IPluginV2DynamicExt* create(const char* op, int32_t normalized, int32_t onesided, int32_t nd)
{
    IPluginCreator* c = g_registry.getPluginCreator(op, "1", "");
    if (!c) return nullptr;
    // like the ONNX parser: walk getFieldNames() and fill the values in that order
    PluginFieldCollection const* names = c->getFieldNames();
    std::vector<PluginField> f;
    int32_t vals[3] = {normalized, onesided, nd};
    for (int32_t i = 0; i < names->nbFields; ++i) {
        const char* n = names->fields[i].name;
        int32_t* v = !strcmp(n, "normalized") ? &vals[0] : !strcmp(n, "onesided") ? &vals[1] : &vals[2];
        f.emplace_back(PluginField{n, v, PluginFieldType::kINT32, 1});
    }
    PluginFieldCollection fc{(int32_t)f.size(), f.data()};
    return static_cast<IPluginV2DynamicExt*>(c->createPlugin(op, &fc));
}

Dims make_dims(int32_t n, const int32_t* d)
{
    Dims r{};
    r.nbDims = n;
    for (int32_t i = 0; i < n; ++i) r.d[i] = d[i];
    return r;
}

bool infer(IPluginV2DynamicExt* p, Dims const& in, Dims* out)
{
    Evaluator ev;
    DimsExprs e{};
    e.nbDims = in.nbDims;
    for (int32_t i = 0; i < in.nbDims; ++i) e.d[i] = ev.constant(in.d[i]);
    DimsExprs o = p->getOutputDimensions(0, &e, 1, ev);
    if (o.nbDims <= 0) return false;
    out->nbDims = o.nbDims;
    for (int32_t i = 0; i < o.nbDims; ++i) out->d[i] = o.d[i]->getConstantValue();
    return true;
}

}  // namespace

extern "C" nvinfer1::IPluginRegistry* getPluginRegistry() noexcept { return &g_registry; }
// Marks this process's "TensorRT" as the test double: a plugin library built against the stand-in header
// registers its creators only when it finds this symbol (see RegisterAtLoad in csrc/plugin/dft_plugins.cpp).
extern "C" int trt_dft_mock_host = 1;

extern "C" {

// also accept creators from a library that was loaded before this one (TRT >= 8.6 loadLibrary path)
int mock_register_from(IPluginCreator* const* (*get_creators)(int32_t&))
{
    int32_t n = 0;
    IPluginCreator* const* cs = get_creators(n);
    int added = 0;
    for (int32_t i = 0; i < n; ++i) added += g_registry.registerCreator(*cs[i], "") ? 1 : 0;
    return added;
}

int mock_registry_count()
{
    int32_t n = 0;
    g_registry.getPluginCreatorList(&n);
    return n;
}

// "name|version|namespace|field:type,field:type,..."
int mock_registry_entry(int i, char* buf, size_t len)
{
    int32_t n = 0;
    IPluginCreator* const* cs = g_registry.getPluginCreatorList(&n);
    if (i < 0 || i >= n) return 1;
    std::string s = std::string(cs[i]->getPluginName()) + "|" + cs[i]->getPluginVersion() + "|" + cs[i]->getPluginNamespace() + "|";
    PluginFieldCollection const* f = cs[i]->getFieldNames();
    for (int32_t j = 0; j < f->nbFields; ++j)
        s += std::string(j ? "," : "") + f->fields[j].name + ":" + std::to_string((int)f->fields[j].type);
    snprintf(buf, len, "%s", s.c_str());
    return 0;
}

// 0 = created, 1 = creator returned nullptr, 2 = no such creator
int mock_try_create(const char* op, int32_t normalized, int32_t onesided, int32_t nd)
{
    if (!g_registry.getPluginCreator(op, "1", "")) return 2;
    IPluginV2DynamicExt* p = create(op, normalized, onesided, nd);
    if (!p) return 1;
    p->destroy();
    return 0;
}

// shape inference; returns nbDims of the output (<= 0 on failure) and fills out_dims
int mock_output_dims(const char* op, int32_t nd, int32_t nb_in, const int32_t* in_dims, int32_t* out_dims)
{
    IPluginV2DynamicExt* p = create(op, 0, 1, nd);
    if (!p) return -1;
    Dims o{};
    const bool ok = infer(p, make_dims(nb_in, in_dims), &o);
    int r = ok ? o.nbDims : -2;
    if (ok) memcpy(out_dims, o.d, sizeof(int32_t) * o.nbDims);
    r = (p->getNbOutputs() == 1) ? r : -3;
    DataType t = DataType::kFLOAT;
    if (p->getOutputDataType(0, &t, 1) != DataType::kFLOAT) r = -4;
    p->destroy();
    return r;
}

// data type of the tensors that mock_workspace_size and mock_lifecycle describe (default fp32)
static DataType g_type = DataType::kFLOAT;
void mock_set_tensor_type(int32_t type) { g_type = (DataType)type; }

// both positions with the given types, as the builder asks once it has settled on a combination
int mock_format_pair_ok(const char* op, int32_t type_in, int32_t type_out)
{
    IPluginV2DynamicExt* p = create(op, 0, 1, 2);
    if (!p) return -1;
    PluginTensorDesc d[2]{};
    for (auto& x : d) { x.format = TensorFormat::kLINEAR; x.scale = 1.f; }
    d[0].type = (DataType)type_in;
    d[1].type = (DataType)type_out;
    const int r = (p->supportsFormatCombination(0, d, 1, 1) && p->supportsFormatCombination(1, d, 1, 1)) ? 1 : 0;
    p->destroy();
    return r;
}

int mock_format_ok(const char* op, int32_t pos, int32_t type, int32_t format)
{
    IPluginV2DynamicExt* p = create(op, 0, 1, 2);
    if (!p) return -1;
    PluginTensorDesc d[2]{};
    for (auto& x : d) { x.type = DataType::kFLOAT; x.format = TensorFormat::kLINEAR; x.scale = 1.f; }
    d[pos].type = (DataType)type;
    d[pos].format = (TensorFormat)format;
    const int r = p->supportsFormatCombination(pos, d, 1, 1) ? 1 : 0;
    p->destroy();
    return r;
}

// serialise a plugin, return its size, copy the bytes, and check type/version/namespace survive a
// deserialise + clone round trip.  Returns the byte count, negative on failure.
int mock_serialize_roundtrip(const char* op, int32_t nd, unsigned char* bytes, size_t cap)
{
    IPluginV2DynamicExt* p = create(op, 0, 1, nd);
    if (!p) return -1;
    p->setPluginNamespace("ns_test");
    const size_t n = p->getSerializationSize();
    if (n > cap) { p->destroy(); return -2; }
    p->serialize(bytes);
    IPluginCreator* c = g_registry.getPluginCreator(op, "1", "");
    IPluginV2* q = c->deserializePlugin(op, bytes, n);
    int r = (int)n;
    if (!q) r = -3;
    else {
        IPluginV2* q2 = q->clone();
        if (!q2 || strcmp(q2->getPluginType(), op) || strcmp(q2->getPluginVersion(), "1")) r = -4;
        std::vector<unsigned char> again(n);
        if (q2) { q2->serialize(again.data()); if (memcmp(again.data(), bytes, n)) r = -5; q2->destroy(); }
        q->destroy();
    }
    IPluginV2* cl = p->clone();
    if (!cl || strcmp(cl->getPluginNamespace(), "ns_test")) r = -6;      // clone keeps the namespace
    if (cl) cl->destroy();
    if (c->deserializePlugin(op, bytes, n - 1) != nullptr) r = -7;        // wrong length must be refused
    p->destroy();
    return r;
}

size_t mock_workspace_size(const char* op, int32_t nd, int32_t nb_in, const int32_t* in_dims)
{
    IPluginV2DynamicExt* p = create(op, 0, 1, nd);
    if (!p) return (size_t)-1;
    Dims in = make_dims(nb_in, in_dims), out{};
    infer(p, in, &out);
    PluginTensorDesc di{in, g_type, TensorFormat::kLINEAR, 1.f}, dox{out, g_type, TensorFormat::kLINEAR, 1.f};
    const size_t w = p->getWorkspaceSize(&di, 1, &dox, 1);
    p->destroy();
    return w;
}

// Full build + inference lifecycle on device buffers.  `skip_configure_on_runtime_clone` exercises
// the case the reference cannot handle (a clone that never saw configurePlugin).
int mock_lifecycle(const char* op, int32_t nd, int32_t nb_in, const int32_t* in_dims, const void* d_in, void* d_out,
                   void* stream, int skip_configure_on_runtime_clone, int32_t* out_dims, int32_t* nb_out)
{
    IPluginV2DynamicExt* built = create(op, 0, 1, nd);
    if (!built) return 1;
    Dims in = make_dims(nb_in, in_dims), out{};
    if (!infer(built, in, &out)) return 2;
    *nb_out = out.nbDims;
    memcpy(out_dims, out.d, sizeof(int32_t) * out.nbDims);
    PluginTensorDesc io[2] = {{in, g_type, TensorFormat::kLINEAR, 1.f}, {out, g_type, TensorFormat::kLINEAR, 1.f}};
    if (!built->supportsFormatCombination(0, io, 1, 1) || !built->supportsFormatCombination(1, io, 1, 1)) return 3;
    // builder: clone, configure, workspace, (timing enqueue skipped), serialize
    IPluginV2DynamicExt* b2 = built->clone();
    DynamicPluginTensorDesc din{io[0], in, in}, dout{io[1], out, out};
    b2->configurePlugin(&din, 1, &dout, 1);
    const size_t ws_bytes = b2->getWorkspaceSize(&io[0], 1, &io[1], 1);
    std::vector<unsigned char> blob(b2->getSerializationSize());
    b2->serialize(blob.data());
    b2->destroy();
    built->destroy();
    // runtime: deserialize, clone per execution context, configure, initialize, enqueue
    IPluginCreator* c = g_registry.getPluginCreator(op, "1", "");
    IPluginV2DynamicExt* rt = static_cast<IPluginV2DynamicExt*>(c->deserializePlugin(op, blob.data(), blob.size()));
    if (!rt) return 4;
    IPluginV2DynamicExt* ctx = rt->clone();
    if (!skip_configure_on_runtime_clone) ctx->configurePlugin(&din, 1, &dout, 1);
    if (ctx->initialize() != 0) return 5;
    void* ws = nullptr;
    if (ws_bytes && cudaMalloc(&ws, ws_bytes) != cudaSuccess) return 6;
    void const* ins[1] = {d_in};
    void* outs[1] = {d_out};
    int rc = ctx->enqueue(&io[0], &io[1], ins, outs, ws, (cudaStream_t)stream);
    if (rc == 0) rc = ctx->enqueue(&io[0], &io[1], ins, outs, ws, (cudaStream_t)stream);   // enqueue is repeatable
    cudaStreamSynchronize((cudaStream_t)stream);
    if (ws) cudaFree(ws);
    ctx->terminate();
    ctx->destroy();
    rt->destroy();
    return rc == 0 ? 0 : 7;
}

// getWorkspaceSize after a dynamic-shape configurePlugin (profile min != max); asked, as TensorRT does, about the
// profile's largest shape.
size_t mock_dynamic_workspace(const char* op, int32_t nd, int32_t nb_in, const int32_t* min_dims, const int32_t* max_dims)
{
    IPluginV2DynamicExt* p = create(op, 0, 1, nd);
    if (!p) return (size_t)-1;
    Dims imin = make_dims(nb_in, min_dims), imax = make_dims(nb_in, max_dims), omin{}, omax{};
    infer(p, imin, &omin);
    infer(p, imax, &omax);
    PluginTensorDesc di{imax, DataType::kFLOAT, TensorFormat::kLINEAR, 1.f}, dox{omax, DataType::kFLOAT, TensorFormat::kLINEAR, 1.f};
    DynamicPluginTensorDesc din{di, imin, imax}, dout{dox, omin, omax};
    p->configurePlugin(&din, 1, &dout, 1);
    const size_t w = p->getWorkspaceSize(&di, 1, &dox, 1);
    p->destroy();
    return w;
}

// What TensorRT does with CUDA graphs: the plugin that was configured (and timed) is one instance, the
// execution context enqueues a clone that never saw configurePlugin, inside a stream capture.  The clone's
// enqueue must then be launch-only (no allocation, no synchronisation), or the capture fails.
// Steps: configure + one ordinary enqueue on the original; clone; capture clone->enqueue; also ask the clone
// for a shape nobody configured while capturing (must return non-zero WITHOUT breaking the capture);
// instantiate, launch the graph `replays` times.  Returns 0, or a step number on failure.
int mock_capture_clone(const char* op, int32_t nd, int32_t nb_in, const int32_t* in_dims, const void* d_in, void* d_out,
                       int32_t replays)
{
    IPluginV2DynamicExt* orig = create(op, 0, 1, nd);
    if (!orig) return 1;
    Dims in = make_dims(nb_in, in_dims), out{};
    if (!infer(orig, in, &out)) return 2;
    PluginTensorDesc io[2] = {{in, DataType::kFLOAT, TensorFormat::kLINEAR, 1.f}, {out, DataType::kFLOAT, TensorFormat::kLINEAR, 1.f}};
    DynamicPluginTensorDesc din{io[0], in, in}, dout{io[1], out, out};
    orig->configurePlugin(&din, 1, &dout, 1);
    const size_t ws_bytes = orig->getWorkspaceSize(&io[0], 1, &io[1], 1);
    void* ws = nullptr;
    if (ws_bytes && cudaMalloc(&ws, ws_bytes) != cudaSuccess) return 3;
    cudaStream_t st = nullptr;
    if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) return 4;
    void const* ins[1] = {d_in};
    void* outs[1] = {d_out};
    int rc = 0;
    if (orig->enqueue(&io[0], &io[1], ins, outs, ws, st) != 0) rc = 5;       // the builder's timing run
    cudaStreamSynchronize(st);
    IPluginV2DynamicExt* ctx = orig->clone();                                // never configured
    if (!ctx) rc = rc ? rc : 6;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    if (!rc && cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) rc = 7;
    if (!rc) {
        if (ctx->enqueue(&io[0], &io[1], ins, outs, ws, st) != 0) rc = 8;
        // a shape no instance has a plan for: must be refused, and the capture must survive it
        int32_t odd[8];
        for (int32_t i = 0; i < nb_in; ++i) odd[i] = in_dims[i];
        odd[0] += 3;
        Dims in2 = make_dims(nb_in, odd), out2{};
        infer(ctx, in2, &out2);
        PluginTensorDesc io2[2] = {{in2, DataType::kFLOAT, TensorFormat::kLINEAR, 1.f}, {out2, DataType::kFLOAT, TensorFormat::kLINEAR, 1.f}};
        if (ctx->enqueue(&io2[0], &io2[1], ins, outs, ws, st) == 0) rc = rc ? rc : 9;
        if (cudaStreamEndCapture(st, &graph) != cudaSuccess || !graph) rc = rc ? rc : 10;
    }
    if (!rc && cudaGraphInstantiate(&exec, graph, 0) != cudaSuccess) rc = 11;
    for (int32_t i = 0; !rc && i < replays; ++i)
        if (cudaGraphLaunch(exec, st) != cudaSuccess) rc = 12;
    if (cudaStreamSynchronize(st) != cudaSuccess) rc = rc ? rc : 13;
    if (exec) cudaGraphExecDestroy(exec);
    if (graph) cudaGraphDestroy(graph);
    cudaGetLastError();
    if (ctx) ctx->destroy();
    orig->destroy();
    cudaStreamDestroy(st);
    if (ws) cudaFree(ws);
    return rc;
}

}  // extern "C"
