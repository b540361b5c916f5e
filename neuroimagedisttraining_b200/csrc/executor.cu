// Native step executor: walks a plan (conv-BN-ReLU-pool blocks + classifier head) and issues every kernel of a forward /
// training step / evaluation batch from native code.  Same launch sequence as ops/engine_b200.py (which keeps a Python path
// for per-kernel profiling); the reference has no counterpart -- its step is autograd over ATen / cuDNN / cuBLAS calls
// (fedml_api/standalone/sailentgrads/my_model_trainer.py:201-235).
#include "common.cuh"
#include "../../include/nidt_b200.h"
#include <mutex>
#include <map>
#include <tuple>
#include <cstddef>

namespace {

struct Box { int bw, bh, bd, bn; };

// box of output positions (bw*bh*bd*bn <= 128) maximising valid rows / (tiles * 128); ties prefer wider boxes
// (the same search as ops/engine_b200.py::choose_box, cached per shape)
Box choose_box(int W, int H, int D, int N) {
    static std::mutex mu;
    static std::map<std::tuple<int, int, int, int>, Box> cache;
    std::lock_guard<std::mutex> lock(mu);
    auto key = std::make_tuple(W, H, D, N);
    auto it = cache.find(key);
    if (it != cache.end()) return it->second;
    Box best{1, 1, 1, 1};
    long best_util = -1;
    int best_tie = 0;
    const double total = (double)W * H * D * N;
    for (int bw = 1; bw <= (W < 128 ? W : 128); ++bw) {
        const int tw = (W + bw - 1) / bw;
        for (int bh = 1; bh <= (H < 128 / bw ? H : 128 / bw); ++bh) {
            const int th = (H + bh - 1) / bh;
            for (int bd = 1; bd <= (D < 128 / (bw * bh) ? D : 128 / (bw * bh)); ++bd) {
                const int td = (D + bd - 1) / bd;
                int bn = 128 / (bw * bh * bd);
                if (bn > N) bn = N;
                if (bn < 1) continue;
                const int tn = (N + bn - 1) / bn;
                const double util = total / ((double)tw * th * td * tn * 128.0);
                const long u = (long)(util * 10000.0 + 0.5);      // round(util, 4)
                const int tie = bw * 1000 + bh;
                if (u > best_util || (u == best_util && tie > best_tie)) { best = Box{bw, bh, bd, bn}; best_util = u; best_tie = tie; }
            }
        }
    }
    cache[key] = best;
    return best;
}

inline float* at(float* slab, int64_t off) { return off >= 0 ? slab + off : nullptr; }
inline const float* at(const float* slab, int64_t off) { return off >= 0 ? slab + off : nullptr; }

#define NIDT_TRY(expr)                 \
    do {                               \
        const int rc_ = (expr);        \
        if (rc_ != 0) return rc_;      \
    } while (0)

struct Ctx {
    const nidt_plan_t* p;
    int B;
    void* st;
    int launches = 0;
    Box cur_box[8];
    const nidt_dyn_t* dyn = nullptr;      // graph capture: step scalars / result pointers come from device memory
    const float* sal_w = nullptr;         // SNIP pass: slab base (weights), score slab and scale for the wgrad saliency epilogues
    float* sal_scores = nullptr;
    float sal_scale = 0.f;
    const float* w_slab = nullptr;        // the slab the forward ran with (the algebraic first-layer gradient reads W and b again)
};

int pack_weights(Ctx& c, const float* slab) {
    const float* w[8];
    void *wf[8], *wd[8];
    int Cout[8], Cin[8], taps[8], n = 0;
    for (int i = 1; i < c.p->n_blocks; ++i, ++n) {
        const nidt_block_t& b = c.p->blocks[i];
        w[n] = at(slab, b.off_w); wf[n] = b.wf; wd[n] = b.wd;
        Cout[n] = b.cout; Cin[n] = b.cin; taps[n] = b.k * b.k * b.k;
    }
    if (n == 0) return 0;
    NIDT_TRY(nidt_pack_conv_weights_multi(n, w, Cout, Cin, taps, wf, wd, c.st));
    ++c.launches;
    return 0;
}

int forward(Ctx& c, const void* x, float* slab, bool train) {
    const nidt_plan_t* p = c.p;
    const int B = c.B;
    c.w_slab = slab;
    const void* prev = nullptr;
    for (int i = 0; i < p->n_blocks; ++i) {
        const nidt_block_t& b = p->blocks[i];
        float* stats = train ? b.stats : nullptr;
        const float* bias = b.has_bias ? at((const float*)slab, b.off_b) : nullptr;
        int tiles = b.stats_tiles;
        // block 0 with nobody reading the full-resolution y afterwards (evaluation; training whose backward works from the pooled tensors):
        // the conv epilogue reduces along W and the pooling kernel finishes from a third of the data
        const bool poolw = i == 0 && b.s2d && b.yw != nullptr && b.aw != nullptr && b.pool_k == 3 && b.pool_s == 3 && b.cout == 64 && b.c_w >= 3 &&
                           (!train || (b.poolw_train && b.bwd0_mode == 2 && b.amax != nullptr && b.ymax != nullptr));
        if (i == 0) {
            if (poolw) NIDT_TRY(nidt_conv_in_s2d_fprop_poolw(x, at(slab, b.off_w), bias, at(slab, b.off_gamma), b.yw, b.aw, stats, B, b.in_d, b.in_h,
                                                             b.in_w, p->input_scale, c.st));
            else if (b.s2d) NIDT_TRY(nidt_conv_in_s2d_fprop(x, at(slab, b.off_w), bias, b.y, stats, B, b.in_d, b.in_h, b.in_w, p->input_scale, c.st));
            else NIDT_TRY(nidt_conv_in_fprop(x, at(slab, b.off_w), bias, b.y, stats, B, b.in_d, b.in_h, b.in_w, b.k, b.stride, b.pad, p->input_scale, c.st));
        } else {
            const Box bx = B == p->max_batch ? Box{b.box[0], b.box[1], b.box[2], b.box[3]} : choose_box(b.c_w, b.c_h, b.c_d, B);
            c.cur_box[i] = bx;
            NIDT_TRY(nidt_conv3d_fprop_tc(prev, b.wf, bias, b.y, stats, B, b.in_d, b.in_h, b.in_w, b.cin, b.c_d, b.c_h, b.c_w, b.cout, b.k, b.stride,
                                          b.pad, bx.bw, bx.bh, bx.bd, bx.bn, b.block_n, c.st));
            if (B != p->max_batch) tiles = nidt_conv3d_m_tiles(B, b.c_d, b.c_h, b.c_w, bx.bw, bx.bh, bx.bd, bx.bn);
        }
        if (train)
            NIDT_TRY(nidt_bn_finalize(b.stats, tiles, b.cout, (double)B * b.c_d * b.c_h * b.c_w, at(slab, b.off_gamma), at(slab, b.off_beta),
                                      at(slab, b.off_rm), at(slab, b.off_rv), at(slab, b.off_nbt), 0.1f, 1e-5f, b.scale, b.shift, b.mean, b.invstd, c.st));
        else
            NIDT_TRY(nidt_bn_eval_prepare(b.cout, at(slab, b.off_gamma), at(slab, b.off_beta), at(slab, b.off_rm), at(slab, b.off_rv), 1e-5f, b.scale,
                                          b.shift, c.st));
        if (poolw)
            NIDT_TRY(nidt_bn_relu_pool3w_fwd(b.yw, b.aw, b.a, train ? b.ymax : nullptr, train ? b.amax : nullptr, B, b.c_d, b.c_h, b.c_w, b.cout, b.scale,
                                             b.shift, b.relu, c.st));
        else if (train && i == 0 && b.bwd0_mode != 0 && b.amax != nullptr && b.ymax != nullptr && b.s2d && b.pool_k == 3 && b.pool_s == 3)
            NIDT_TRY(nidt_bn_relu_pool3_fwd_amax(b.y, b.a, b.ymax, b.amax, B, b.c_d, b.c_h, b.c_w, b.cout, b.scale, b.shift, b.relu, c.st));
        else
            NIDT_TRY(nidt_bn_relu_pool_fwd(b.y, b.a, (train && b.ymax != nullptr) ? b.ymax : nullptr, B, b.c_d, b.c_h, b.c_w, b.cout, b.pool_k, b.pool_s,
                                           b.scale, b.shift, b.relu, c.st));
        c.launches += 3;
        prev = b.a;
    }
    return 0;
}

int head(Ctx& c, const float* y, const float* slab, float* grad, bool train, double* counters, int double_sigmoid, uint64_t seed) {
    const nidt_plan_t* p = c.p;
    const nidt_block_t& last = p->blocks[p->n_blocks - 1];
    const bool g = train && grad != nullptr;
    if (c.dyn != nullptr) {
        NIDT_TRY(nidt_head_fwd_bwd_dyn(last.a, y, at(slab, p->off_fc1_w), at(slab, p->off_fc1_b), at(slab, p->off_fc2_w), at(slab, p->off_fc2_b),
                                       g ? at(grad, p->off_fc1_w) : nullptr, g ? at(grad, p->off_fc1_b) : nullptr, g ? at(grad, p->off_fc2_w) : nullptr,
                                       g ? at(grad, p->off_fc2_b) : nullptr, train ? last.da : nullptr, p->eval_logits, c.B, p->head_s, p->head_c,
                                       p->head_hidden, p->p_drop, train ? 1 : 0, double_sigmoid, c.dyn, c.st));
        ++c.launches;
        return 0;
    }
    NIDT_TRY(nidt_head_fwd_bwd(last.a, y, at(slab, p->off_fc1_w), at(slab, p->off_fc1_b), at(slab, p->off_fc2_w), at(slab, p->off_fc2_b),
                               g ? at(grad, p->off_fc1_w) : nullptr, g ? at(grad, p->off_fc1_b) : nullptr, g ? at(grad, p->off_fc2_w) : nullptr,
                               g ? at(grad, p->off_fc2_b) : nullptr, train ? last.da : nullptr, train ? p->loss_acc : nullptr, counters, p->eval_logits,
                               c.B, p->head_s, p->head_c, p->head_hidden, p->p_drop, seed, train ? 1 : 0, double_sigmoid, c.st));
    ++c.launches;
    return 0;
}

int backward(Ctx& c, const void* x, float* grad) {
    const nidt_plan_t* p = c.p;
    const int B = c.B;
    for (int i = p->n_blocks - 1; i >= 0; --i) {
        const nidt_block_t& b = p->blocks[i];
        if (i == 0 && b.bwd0_mode == 2 && b.amax != nullptr && b.ymax != nullptr && b.gram != nullptr && b.alg_scratch != nullptr && b.s2d &&
            b.pool_k == 3 && b.pool_s == 3 && b.cout == 64) {
            // first block, algebraic form: BatchNorm-backward reductions over the pooled tensors, then the weight gradient from one
            // input patch per pooling window + the cached Gram matrix of the batch's input patches; y and dY are not touched
            NIDT_TRY(nidt_bn_relu_pool3_bwd_stats(b.ymax, b.da, B, b.c_d, b.c_h, b.c_w, b.cout, b.scale, b.shift, b.mean, b.invstd, b.bwd_partials,
                                                  b.c12, at(grad, b.off_gamma), at(grad, b.off_beta), b.relu, c.st));
            const bool sal = c.sal_scores != nullptr;
            NIDT_TRY(nidt_conv_in_alg_wgrad(x, b.amax, b.da, at(c.w_slab, b.off_w), b.has_bias ? at(c.w_slab, b.off_b) : nullptr, b.scale, b.mean,
                                            b.invstd, b.c12, b.gram, at(grad, b.off_w), b.has_bias ? at(grad, b.off_b) : nullptr, b.alg_scratch, B,
                                            b.in_d, b.in_h, b.in_w, p->input_scale, sal ? at(c.sal_w, b.off_w) : nullptr,
                                            sal ? at(c.sal_scores, b.off_w) : nullptr, c.sal_scale, c.st));
            c.launches += 4;
            continue;
        }
        if (i == 0 && b.bwd0_mode == 1 && b.amax != nullptr && b.ymax != nullptr && b.s2d && b.pool_k == 3 && b.pool_s == 3) {
            // first block: only the BatchNorm-backward reductions run as kernels of their own; dY is formed inside the weight-gradient
            // kernel from y, the winning taps and da (never written to HBM)
            float* gb = b.has_bias ? at(grad, b.off_b) : nullptr;
            NIDT_TRY(nidt_bn_relu_pool3_bwd_stats(b.ymax, b.da, B, b.c_d, b.c_h, b.c_w, b.cout, b.scale, b.shift, b.mean, b.invstd, b.bwd_partials,
                                                  b.c12, at(grad, b.off_gamma), at(grad, b.off_beta), b.relu, c.st));
            const bool sal = c.sal_scores != nullptr;
            NIDT_TRY(nidt_conv_in_s2d_wgrad_fused(x, b.y, b.amax, b.da, b.scale, b.mean, b.invstd, b.c12, at(grad, b.off_w), gb, b.wg_scratch, B,
                                                  b.in_d, b.in_h, b.in_w, p->input_scale, sal ? at(c.sal_w, b.off_w) : nullptr,
                                                  sal ? at(c.sal_scores, b.off_w) : nullptr, c.sal_scale, c.st));
            c.launches += 4;
            continue;
        }
        NIDT_TRY(nidt_bn_relu_pool_bwd(b.y, b.ymax, b.da, b.dy, B, b.c_d, b.c_h, b.c_w, b.cout, b.pool_k, b.pool_s, b.scale, b.shift, b.mean, b.invstd,
                                       b.bwd_partials, b.c12, at(grad, b.off_gamma), at(grad, b.off_beta), b.relu, c.st));
        c.launches += 3;
        if (i == 0) {
            float* gb = b.has_bias ? at(grad, b.off_b) : nullptr;
            if (b.s2d && c.sal_scores != nullptr)
                NIDT_TRY(nidt_conv_in_s2d_wgrad_saliency(x, b.dy, at(grad, b.off_w), gb, b.wg_scratch, B, b.in_d, b.in_h, b.in_w, p->input_scale,
                                                         at(c.sal_w, b.off_w), at(c.sal_scores, b.off_w), c.sal_scale, c.st));
            else if (b.s2d) NIDT_TRY(nidt_conv_in_s2d_wgrad(x, b.dy, at(grad, b.off_w), gb, b.wg_scratch, B, b.in_d, b.in_h, b.in_w, p->input_scale, c.st));
            else NIDT_TRY(nidt_conv_in_wgrad(x, b.dy, at(grad, b.off_w), gb, b.wg_scratch, B, b.in_d, b.in_h, b.in_w, b.k, b.stride, b.pad, p->input_scale, c.st));
            c.launches += 2;
        } else {
            const nidt_block_t& prev = p->blocks[i - 1];
            const Box bx = c.cur_box[i];
            int splits = b.splits;
            if (B != p->max_batch) {
                const int mt = nidt_conv3d_m_tiles(B, b.c_d, b.c_h, b.c_w, bx.bw, bx.bh, bx.bd, bx.bn);
                if (splits > mt) splits = mt;
                if (splits < 1) splits = 1;
            }
            if (c.sal_scores != nullptr)
                NIDT_TRY(nidt_conv3d_wgrad_tc_saliency(b.dy, prev.a, at(grad, b.off_w), b.wg_scratch, B, b.in_d, b.in_h, b.in_w, b.cin, b.c_d, b.c_h,
                                                       b.c_w, b.cout, b.k, b.stride, b.pad, bx.bw, bx.bh, bx.bd, bx.bn, splits, at(c.sal_w, b.off_w),
                                                       at(c.sal_scores, b.off_w), c.sal_scale, c.st));
            else
                NIDT_TRY(nidt_conv3d_wgrad_tc(b.dy, prev.a, at(grad, b.off_w), b.wg_scratch, B, b.in_d, b.in_h, b.in_w, b.cin, b.c_d, b.c_h, b.c_w, b.cout,
                                              b.k, b.stride, b.pad, bx.bw, bx.bh, bx.bd, bx.bn, splits, c.st));
            // dgrad = fprop of dY with the transposed, tap-flipped weights and pad' = k-1-pad
            const Box db = B == p->max_batch ? Box{b.dbox[0], b.dbox[1], b.dbox[2], b.dbox[3]} : choose_box(b.in_w, b.in_h, b.in_d, B);
            NIDT_TRY(nidt_conv3d_fprop_tc(b.dy, b.wd, nullptr, prev.da, nullptr, B, b.c_d, b.c_h, b.c_w, b.cout, b.in_d, b.in_h, b.in_w, b.cin, b.k, 1,
                                          b.k - 1 - b.pad, db.bw, db.bh, db.bd, db.bn, b.dgrad_block_n, c.st));
            c.launches += 3;
        }
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------ CUDA-graph replays
struct GraphKey {
    const void* plan; const void* x; const void* y; const void* slab; const void* grad; const void* mom; const void* mask; const void* dyn;
    int64_t n_params; int batch, kind, flag, device;
    bool operator<(const GraphKey& o) const {
        return std::tie(plan, x, y, slab, grad, mom, mask, dyn, n_params, batch, kind, flag, device) <
               std::tie(o.plan, o.x, o.y, o.slab, o.grad, o.mom, o.mask, o.dyn, o.n_params, o.batch, o.kind, o.flag, o.device);
    }
};
struct GraphVal { cudaGraphExec_t exec; int launches; };
std::mutex g_graph_mu;
std::map<GraphKey, GraphVal> g_graphs;
std::map<int, cudaStream_t> g_capture_streams;        // per device: capture needs a non-default stream
constexpr size_t kMaxGraphs = 1024;

__global__ void set_dyn_kernel(nidt_dyn_t* dyn, const nidt_dyn_t v) { *dyn = v; }

void drop_graphs(const void* plan) {
    for (auto it = g_graphs.begin(); it != g_graphs.end();) {
        if (plan == nullptr || it->first.plan == plan) { cudaGraphExecDestroy(it->second.exec); it = g_graphs.erase(it); }
        else ++it;
    }
}

// Records `body` (which launches on the given stream) into an executable graph.
template <class Body>
int capture(int device, GraphVal* out, Body body) {
    cudaStream_t& cap = g_capture_streams[device];
    if (cap == nullptr) NIDT_CUDA_CHECK(cudaStreamCreateWithFlags(&cap, cudaStreamNonBlocking));
    NIDT_CUDA_CHECK(cudaStreamBeginCapture(cap, cudaStreamCaptureModeRelaxed));
    int launches = 0;
    const int rc = body((void*)cap, &launches);
    cudaGraph_t graph = nullptr;
    const cudaError_t e = cudaStreamEndCapture(cap, &graph);
    if (rc != 0) { if (graph) cudaGraphDestroy(graph); return rc; }
    if (e != cudaSuccess || graph == nullptr) { nidt_set_error("graph capture failed: %s", cudaGetErrorString(e)); cudaGetLastError(); return 1; }
    cudaGraphExec_t exec = nullptr;
    const cudaError_t ei = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ei != cudaSuccess) { nidt_set_error("cudaGraphInstantiate: %s", cudaGetErrorString(ei)); cudaGetLastError(); return 1; }
    out->exec = exec;
    out->launches = launches;
    return 0;
}

template <class Body>
int replay(const GraphKey& key, const nidt_dyn_t& values, nidt_dyn_t* dyn, int* launches, void* stream, Body body) {
    std::lock_guard<std::mutex> lock(g_graph_mu);
    auto it = g_graphs.find(key);
    if (it == g_graphs.end()) {
        if (g_graphs.size() >= kMaxGraphs) drop_graphs(nullptr);
        GraphVal v{};
        NIDT_TRY(capture(key.device, &v, body));
        it = g_graphs.emplace(key, v).first;
    }
    set_dyn_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(dyn, values);
    NIDT_LAUNCH_CHECK();
    NIDT_CUDA_CHECK(cudaGraphLaunch(it->second.exec, (cudaStream_t)stream));
    if (launches != nullptr) *launches = it->second.launches + 1;
    return 0;
}

}  // namespace

NIDT_API int nidt_plan_train_step_graph(const nidt_plan_t* plan, const void* x_u8, const float* y, int batch, float* slab, float* grad, float* mom,
                                        const float* mask, int64_t n_params, float lr, float momentum, float wd, float clip_norm, int first_step,
                                        uint64_t seed, int do_update, float* loss_acc, nidt_dyn_t* dyn, int* launches, void* stream) {
    NIDT_REQUIRE(plan != nullptr && plan->n_blocks >= 1 && plan->n_blocks <= 8, "nidt_plan: bad plan");
    NIDT_REQUIRE(batch >= 1 && batch <= plan->max_batch, "nidt_plan_train_step_graph: batch %d exceeds the plan's %d", batch, plan->max_batch);
    NIDT_REQUIRE(dyn != nullptr, "nidt_plan_train_step_graph: dyn is null");
    int device = 0;
    NIDT_CUDA_CHECK(cudaGetDevice(&device));
    float* mom_used = (do_update && momentum != 0.f) ? mom : nullptr;
    const GraphKey key{plan, x_u8, y, slab, grad, mom_used, do_update ? mask : nullptr, dyn, n_params, batch, 0,
                       do_update | (plan->blocks[0].bwd0_mode << 1), device};
    nidt_dyn_t v{lr, momentum, wd, clip_norm, first_step, 0, seed, loss_acc, nullptr};
    const nidt_plan_t snap = *plan;       // the launch sequence bakes the plan's buffers in by value
    return replay(key, v, dyn, launches, stream, [&](void* st, int* n) -> int {
        Ctx c{&snap, batch, st};
        c.dyn = dyn;
        NIDT_TRY(forward(c, x_u8, slab, true));
        NIDT_TRY(head(c, y, slab, grad, true, nullptr, 1, 0));
        NIDT_TRY(backward(c, x_u8, grad));
        if (do_update) {
            NIDT_TRY(nidt_masked_sgd_step_dyn(slab, grad, mom_used, mask, n_params, snap.norm_scratch, dyn, st));
            c.launches += 2;
            NIDT_TRY(pack_weights(c, slab));
        }
        *n = c.launches;
        return 0;
    });
}

NIDT_API int nidt_plan_eval_graph(const nidt_plan_t* plan, const void* x_u8, const float* y, int batch, const float* slab, double* counters,
                                  int double_sigmoid, nidt_dyn_t* dyn, int* launches, void* stream) {
    NIDT_REQUIRE(plan != nullptr && plan->n_blocks >= 1 && plan->n_blocks <= 8, "nidt_plan: bad plan");
    NIDT_REQUIRE(batch >= 1 && batch <= plan->max_batch, "nidt_plan_eval_graph: batch %d exceeds the plan's %d", batch, plan->max_batch);
    NIDT_REQUIRE(dyn != nullptr && counters != nullptr, "nidt_plan_eval_graph: dyn / counters is null");
    int device = 0;
    NIDT_CUDA_CHECK(cudaGetDevice(&device));
    const GraphKey key{plan, x_u8, y, slab, nullptr, nullptr, nullptr, dyn, 0, batch, 1, double_sigmoid, device};
    nidt_dyn_t v{0.f, 0.f, 0.f, 0.f, 0, 0, 0, nullptr, counters};
    const nidt_plan_t snap = *plan;
    return replay(key, v, dyn, launches, stream, [&](void* st, int* n) -> int {
        Ctx c{&snap, batch, st};
        c.dyn = dyn;
        NIDT_TRY(forward(c, x_u8, const_cast<float*>(slab), false));
        NIDT_TRY(head(c, y, slab, nullptr, false, nullptr, double_sigmoid, 0));
        *n = c.launches;
        return 0;
    });
}

NIDT_API int nidt_plan_graph_clear(const nidt_plan_t* plan) {
    std::lock_guard<std::mutex> lock(g_graph_mu);
    drop_graphs(plan);
    return 0;
}
NIDT_API int nidt_plan_graph_count() {
    std::lock_guard<std::mutex> lock(g_graph_mu);
    return (int)g_graphs.size();
}

NIDT_API int nidt_plan_pack_weights(const nidt_plan_t* plan, const float* slab, void* stream) {
    NIDT_REQUIRE(plan != nullptr && plan->n_blocks >= 1 && plan->n_blocks <= 8, "nidt_plan: bad plan");
    Ctx c{plan, plan->max_batch, stream};
    return pack_weights(c, slab);
}

NIDT_API int nidt_plan_train_step(const nidt_plan_t* plan, const void* x_u8, const float* y, int batch, float* slab, float* grad, float* mom,
                                  const float* mask, int64_t n_params, float lr, float momentum, float wd, float clip_norm, int first_step,
                                  uint64_t seed, int do_update, int* launches, void* stream) {
    NIDT_REQUIRE(plan != nullptr && plan->n_blocks >= 1 && plan->n_blocks <= 8, "nidt_plan: bad plan");
    NIDT_REQUIRE(batch >= 1 && batch <= plan->max_batch, "nidt_plan_train_step: batch %d exceeds the plan's %d", batch, plan->max_batch);
    Ctx c{plan, batch, stream};
    NIDT_TRY(forward(c, x_u8, slab, true));
    NIDT_TRY(head(c, y, slab, grad, true, nullptr, 1, seed));
    NIDT_TRY(backward(c, x_u8, grad));
    if (do_update) {
        NIDT_TRY(nidt_masked_sgd_step(slab, grad, momentum != 0.f ? mom : nullptr, mask, n_params, lr, momentum, wd, clip_norm, first_step,
                                      plan->norm_scratch, nullptr, stream));
        c.launches += 2;
        NIDT_TRY(pack_weights(c, slab));
    }
    if (launches != nullptr) *launches = c.launches;
    return 0;
}

NIDT_API int nidt_plan_saliency_step(const nidt_plan_t* plan, const void* x_u8, const float* y, int batch, float* slab, float* grad, float* scores,
                                     float scale, int64_t n_params, uint64_t seed, int* launches, void* stream) {
    NIDT_REQUIRE(plan != nullptr && plan->n_blocks >= 1 && plan->n_blocks <= 8, "nidt_plan: bad plan");
    NIDT_REQUIRE(batch >= 1 && batch <= plan->max_batch, "nidt_plan_saliency_step: batch %d exceeds the plan's %d", batch, plan->max_batch);
    NIDT_REQUIRE(scores != nullptr && plan->blocks[0].s2d, "nidt_plan_saliency_step: needs a score slab and the space-to-depth first layer");
    Ctx c{plan, batch, stream};
    c.sal_w = slab; c.sal_scores = scores; c.sal_scale = scale;
    NIDT_TRY(forward(c, x_u8, slab, true));
    NIDT_TRY(head(c, y, slab, grad, true, nullptr, 1, seed));
    NIDT_TRY(backward(c, x_u8, grad));
    // the classifier (everything from fc1.weight to the end of the parameters): one flat pass over that segment
    const int64_t o = plan->off_fc1_w;
    NIDT_REQUIRE(o >= 0 && o <= n_params, "nidt_plan_saliency_step: bad classifier offset");
    if (n_params > o) {
        NIDT_TRY(nidt_saliency_accumulate(scores + o, grad + o, slab + o, n_params - o, scale, stream));
        ++c.launches;
    }
    if (launches != nullptr) *launches = c.launches;
    return 0;
}

NIDT_API int nidt_plan_eval(const nidt_plan_t* plan, const void* x_u8, const float* y, int batch, const float* slab, double* counters,
                            int double_sigmoid, int* launches, void* stream) {
    NIDT_REQUIRE(plan != nullptr && plan->n_blocks >= 1 && plan->n_blocks <= 8, "nidt_plan: bad plan");
    NIDT_REQUIRE(batch >= 1 && batch <= plan->max_batch, "nidt_plan_eval: batch %d exceeds the plan's %d", batch, plan->max_batch);
    Ctx c{plan, batch, stream};
    NIDT_TRY(forward(c, x_u8, const_cast<float*>(slab), false));
    NIDT_TRY(head(c, y, slab, nullptr, false, counters, double_sigmoid, 0));
    if (launches != nullptr) *launches = c.launches;
    return 0;
}

// layout probes for foreign-language bindings (tests compare them with the ctypes mirror in ops/engine_b200.py)
NIDT_API int nidt_plan_sizeof() { return (int)sizeof(nidt_plan_t); }
NIDT_API int nidt_block_sizeof() { return (int)sizeof(nidt_block_t); }
NIDT_API int nidt_plan_offsetof(int field) {
    switch (field) {
        case 0: return (int)offsetof(nidt_plan_t, blocks);
        case 1: return (int)offsetof(nidt_plan_t, off_fc1_w);
        case 2: return (int)offsetof(nidt_plan_t, head_s);
        case 3: return (int)offsetof(nidt_plan_t, p_drop);
        case 4: return (int)offsetof(nidt_plan_t, loss_acc);
        case 5: return (int)offsetof(nidt_plan_t, norm_scratch);
        case 10: return (int)offsetof(nidt_block_t, off_w);
        case 11: return (int)offsetof(nidt_block_t, y);
        case 12: return (int)offsetof(nidt_block_t, wf);
        case 13: return (int)offsetof(nidt_block_t, stats_tiles);
        case 14: return (int)offsetof(nidt_block_t, box);
        case 15: return (int)offsetof(nidt_block_t, dbox);
        default: return -1;
    }
}

// the M-tile box the executor picks for a partial batch (same search as ops/engine_b200.py::choose_box); out = {bw, bh, bd, bn}
NIDT_API int nidt_choose_box(int W, int H, int D, int N, int* out) {
    NIDT_REQUIRE(W >= 1 && H >= 1 && D >= 1 && N >= 1 && out != nullptr, "nidt_choose_box: bad arguments");
    const Box b = choose_box(W, H, D, N);
    out[0] = b.bw; out[1] = b.bh; out[2] = b.bd; out[3] = b.bn;
    return 0;
}
