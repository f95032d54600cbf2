// backend_graph.cu -- graph_compute of the B200 ggml backend: plan -> fuse -> (capture once) -> replay.
//
// Replaces ggml_backend_cuda_graph_compute (ggml-cuda.cu:2508-2823).  The reference walks the node list and launches one
// kernel chain per node (plus a quantize_q8_1 launch per matmul), and patches the KV-store cpy nodes of its CUDA graph with
// cudaGraphExecKernelNodeSetParams every token (:2593-2650).  Here:
//   * plan:   the node list is lowered to a list of Steps.  Adjacent RMS_NORM -> MUL and SILU -> MUL pairs become one kernel;
//             decode-regime MUL_MATs that read the same activation (Q/K/V, gate/up) are hoisted into one multi-job GEMV launch
//             when no node in between touches the hoisted node's output bytes.
//   * capture: the Step list is captured into one CUDA graph keyed by a hash over everything that is baked into kernel
//             parameters (ops, shapes, strides, op_params, data pointers) EXCEPT the destination of CPY nodes: those read
//             their base pointer from a device slot table that is refreshed by one small H2D copy before each replay, so the
//             moving KV-cache write position does not invalidate the graph.
//   * replay: cudaGraphLaunch.
// There is no CPU fallback: a node supports_op said no to never reaches this file, and anything else aborts like the reference.
#include <algorithm>
#include <unordered_map>
#include <unordered_set>

#include "../../include/b200_comm.h"
#include "../../include/b200_ggml_ops.h"
#include <chrono>

#include "backend_exec.h"

namespace b200 {

// Sibling GEMVs of two K-quant weight types in one launch (attn_q / attn_k Q4_K + attn_v Q6_K): built and bit-exact
// (tests/test_gpu_kernels.py::test_gemv_two_weight_types_in_one_launch), but measured neutral on the 7B step (211 -> 195 launches,
// 507.8 -> 504.9 tok/s: the merged launch needs the larger Q6_K stage, so 2 ring stages instead of 3, and under PDL the saved
// launch boundary was mostly hidden already) and it lowers the GEMV family's bytes/s; therefore opt-in (B200_MIXED_TYPES=1).
static const bool g_mixed_types = getenv("B200_MIXED_TYPES") != nullptr;
static inline bool rows_are_contiguous(const ggml_tensor *w) { return w->nb[1] == row_size(w->type, w->ne[0]); }

enum StepKind { STEP_NODE, STEP_RMS_MUL, STEP_SILU_MUL, STEP_GEMV, STEP_ROPE_KV, STEP_AR_ADD, STEP_MOE_ROUTER };

struct Step {
    int kind;
    const ggml_tensor *n;               // the node (STEP_NODE), or the first node of the fused group
    const ggml_tensor *n2;              // second node of a fused pair
    std::vector<const ggml_tensor *> group;  // STEP_GEMV members; STEP_ROPE_KV: {CPY K, CPY V}
    std::vector<const ggml_tensor *> adds;   // STEP_GEMV: per member the ADD node folded into its epilogue (or NULL)
    const ggml_tensor *silu_mul = nullptr;   // STEP_GEMV of (gate, up): the MUL node of SILU(gate) * up folded into the epilogue
    int gate_idx = 0;                   // ... and which member of `group` is the gate
    int slot = -1, slot2 = -1;          // CPY nodes: index into the slot table, else -1
    bool store_norm = false, store_norm_mul = false;  // STEP_GEMV with a folded norm: this launch also stores the RMS_NORM / MUL values
    Step() : kind(STEP_NODE), n(nullptr), n2(nullptr) {}
};

struct CachedGraph {
    cudaGraphExec_t exec = nullptr;
    int64_t launches = 0;  // kernels inside one replay
    int n_slots = 0;
    uint64_t last_used = 0;
    uint64_t check = 0;    // second, independent hash of the graph: compared on every cache hit
};

struct GraphExec {
    int device;
    cudaStream_t stream;
    void *scratch = nullptr;  size_t scratch_bytes = 0;
    void *fa_ws = nullptr;    size_t fa_ws_bytes = 0;
    void **slots_dev = nullptr;  int slots_cap = 0;
    std::vector<void *> slots_host;
    std::unordered_map<uint64_t, CachedGraph> cache;
    uint64_t tick = 0;
    std::vector<Step> steps;  // plan of the graph being computed (reused storage)
};

GraphExec *graph_exec_new(int device) {
    GraphExec *g = new GraphExec();
    g->device = device;
    B200_CHECK_CUDA(cudaSetDevice(device));
    B200_CHECK_CUDA(cudaStreamCreateWithFlags(&g->stream, cudaStreamNonBlocking));
    return g;
}

static void drop_cache(GraphExec *g) {
    for (auto &kv : g->cache) if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    g->cache.clear();
}

void graph_exec_free(GraphExec *g) {
    if (!g) return;
    cudaSetDevice(g->device);
    cudaStreamSynchronize(g->stream);
    drop_cache(g);
    if (g->scratch) cudaFree(g->scratch);
    if (g->fa_ws) cudaFree(g->fa_ws);
    if (g->slots_dev) cudaFree(g->slots_dev);
    cudaStreamDestroy(g->stream);
    delete g;
}

cudaStream_t graph_exec_stream(GraphExec *g) { return g->stream; }

// ------------------------------------------------------------------------------------------------ planning
static bool is_view_op(int op) {
    return op == GGML_OP_NONE || op == GGML_OP_RESHAPE || op == GGML_OP_VIEW || op == GGML_OP_PERMUTE || op == GGML_OP_TRANSPOSE;
}
static bool is_cpy_op(int op) { return op == GGML_OP_CPY || op == GGML_OP_DUP || op == GGML_OP_CONT; }

static bool ranges_overlap(const void *a, size_t na, const void *b, size_t nb_) {
    const char *pa = (const char *)a, *pb = (const char *)b;
    return pa < pb + nb_ && pb < pa + na;
}
// does node `m` read or write any byte of [p, p+len)?
static bool node_touches(const ggml_tensor *m, const void *p, size_t len) {
    if (m->data && ranges_overlap(m->data, nbytes(m), p, len)) return true;
    for (int i = 0; i < GGML_MAX_SRC; i++) {
        const ggml_tensor *s = m->src[i];
        if (s && s->data && ranges_overlap(s->data, nbytes(s), p, len)) return true;
    }
    return false;
}

static bool rms_mul_pair(const ggml_tensor *a, const ggml_tensor *b) {
    if (a->op != GGML_OP_RMS_NORM || b->op != GGML_OP_MUL) return false;
    const ggml_tensor *w = b->src[0] == a ? b->src[1] : (b->src[1] == a ? b->src[0] : nullptr);
    if (!w || w->type != GGML_TYPE_F32 || !is_contiguous(w)) return false;
    if (w->ne[0] != a->ne[0] || w->ne[1] != 1 || w->ne[2] != 1 || w->ne[3] != 1) return false;
    return same_shape(a, b) && is_contiguous(a) && b->type == GGML_TYPE_F32 && b->nb[0] == 4 && a->src[0]->nb[0] == 4;
}
static bool silu_mul_pair(const ggml_tensor *a, const ggml_tensor *b) {
    if (a->op != GGML_OP_UNARY || a->op_params[0] != GGML_UNARY_OP_SILU || b->op != GGML_OP_MUL) return false;
    const ggml_tensor *u = b->src[0] == a ? b->src[1] : (b->src[1] == a ? b->src[0] : nullptr);
    if (!u || u->type != GGML_TYPE_F32) return false;
    return same_shape(a, b) && same_shape(a, u) && is_contiguous(a) && is_contiguous(b) && is_contiguous(u) && is_contiguous(a->src[0]);
}

// ---- graph facts the fusion rules need
struct GraphInfo {
    const ggml_cgraph *cg;
    std::unordered_map<const ggml_tensor *, int> index;     // node -> position
    std::unordered_map<const ggml_tensor *, int> n_uses;    // tensor -> number of nodes that have it as a src
    std::unordered_map<const ggml_tensor *, const ggml_tensor *> last_user;
    explicit GraphInfo(const ggml_cgraph *g) : cg(g) {
        for (int i = 0; i < g->n_nodes; i++) {
            const ggml_tensor *t = g->nodes[i];
            index[t] = i;
            for (int k = 0; k < GGML_MAX_SRC; k++) {
                if (t->src[k]) { n_uses[t->src[k]]++; last_user[t->src[k]] = t; }
            }
        }
    }
    int pos(const ggml_tensor *t) const { auto it = index.find(t); return it == index.end() ? -1 : it->second; }
    int uses(const ggml_tensor *t) const { auto it = n_uses.find(t); return it == n_uses.end() ? 0 : it->second; }
    // t is read by exactly one node, and nothing outside the graph can want it (outputs are flagged or named result*)
    const ggml_tensor *sole_user(const ggml_tensor *t) const {
        if (uses(t) != 1 || (t->flags & GGML_TENSOR_FLAG_OUTPUT) || strncmp(t->name, "result", 6) == 0) return nullptr;
        if (pos(t) == cg->n_nodes - 1) return nullptr;
        return last_user.find(t)->second;
    }
};

static bool same_op_params(const ggml_tensor *a, const ggml_tensor *b) { return memcmp(a->op_params, b->op_params, sizeof(a->op_params)) == 0; }

// May the nodes `members` (positions ascending) all execute at position `at` (>= first member, <= last member)?
//   * a member moved LATER (pos < at): no outside node in (pos, at] reads or writes its output, or writes one of its inputs;
//   * a member moved EARLIER (pos > at): its inputs exist at `at` (checked by the caller), and no outside node in [at, pos)
//     reads or writes its output bytes.
//   * `done` nodes were absorbed by an earlier step and normally ran at or before that step's position -- except the members of a
//     deferred step (ROPE + KV store, emitted at its LAST member): exec_pos[i] != i says where node i really runs.  A member whose
//     order against such a node changes is checked against it like against any outside node.
static bool can_gather(const GraphInfo &gi, const std::vector<int> &members, int at, const std::vector<char> &done, const std::vector<int> &exec_pos) {
    auto is_member = [&](int i) { for (int m : members) if (m == i) return true; return false; };
    for (int m : members) {
        const ggml_tensor *t = gi.cg->nodes[m];
        const size_t len = nbytes(t);
        for (int i = 0; i < (int)exec_pos.size(); i++) {
            const int e = exec_pos[i];
            if (e == i || is_member(i) || is_member(e)) continue;
            if ((i < m) == (e < at)) continue;  // same order as in the graph
            const ggml_tensor *o = gi.cg->nodes[i];
            if (is_view_op(o->op)) continue;
            if (node_touches(o, t->data, len)) return false;
            if (o->data) {
                for (int k = 0; k < GGML_MAX_SRC; k++) {
                    const ggml_tensor *sk = t->src[k];
                    if (sk && sk->data && ranges_overlap(o->data, nbytes(o), sk->data, nbytes(sk))) return false;
                }
            }
        }
        const int lo = m < at ? m + 1 : at, hi = m < at ? at : m - 1;
        for (int i = lo; i <= hi; i++) {
            if (is_member(i) || (done[i] && i != at)) continue;
            const ggml_tensor *o = gi.cg->nodes[i];
            if (is_view_op(o->op)) continue;
            if (node_touches(o, t->data, len)) return false;
            if (m < at && o->data) {
                for (int k = 0; k < GGML_MAX_SRC; k++) {
                    const ggml_tensor *sk = t->src[k];
                    if (sk && sk->data && ranges_overlap(o->data, nbytes(o), sk->data, nbytes(sk))) return false;
                }
            }
        }
    }
    return true;
}

// ROPE(Q) at position i: find ROPE(K), CPY(K -> cache), CPY(V -> cache) behind it (llm_build_kv order) -> one launch at the last one
static bool match_rope_kv(const GraphInfo &gi, int i, const std::vector<char> &done, int out[4]) {
    const ggml_cgraph *cg = gi.cg;
    const ggml_tensor *rq = cg->nodes[i];
    if (rq->op != GGML_OP_ROPE || (rq->op_params[2] & GGML_ROPE_TYPE_NEOX) || rq->type != GGML_TYPE_F32) return false;
    if (!is_contiguous(rq) || !is_contiguous(rq->src[0]) || rq->ne[3] != 1 || rq->ne[0] % 32) return false;
    // the K-shift graph (build_k_shift, src/llama.cpp:10144-10192) is a chain of per-layer ROPEs over cache views / casts with CPYs
    // back into the cache: not this pattern.  Q and K here are fresh projections, V is never a ROPE result.
    {
        const ggml_tensor *q0 = rq->src[0];
        while (q0 && is_view_op(q0->op) && q0->op != GGML_OP_NONE && q0->src[0]) q0 = q0->src[0];
        if (!q0 || is_cpy_op(q0->op) || q0->op == GGML_OP_NONE) return false;
    }
    int rk = -1, ck = -1, cv = -1;
    const int horizon = i + 24 < cg->n_nodes ? i + 24 : cg->n_nodes;
    for (int j = i + 1; j < horizon; j++) {
        const ggml_tensor *t = cg->nodes[j];
        if (done[j]) continue;
        if (rk < 0) {
            if (t->op == GGML_OP_ROPE && t->src[1] == rq->src[1] && t->src[2] == rq->src[2] && same_op_params(t, rq) && t->ne[0] == rq->ne[0] &&
                t->ne[2] == rq->ne[2] && t->ne[3] == 1 && t->type == GGML_TYPE_F32 && is_contiguous(t) && is_contiguous(t->src[0])) rk = j;
            else if (t->op == GGML_OP_ROPE || t->op == GGML_OP_FLASH_ATTN_EXT) return false;
            continue;
        }
        if (t->op != GGML_OP_CPY) { if (t->op == GGML_OP_FLASH_ATTN_EXT || t->op == GGML_OP_ROPE) break; continue; }
        const ggml_tensor *src = t->src[0];
        if (t->type != GGML_TYPE_F16 && t->type != GGML_TYPE_Q8_0) continue;
        if (!is_contiguous(t) || src->type != GGML_TYPE_F32 || !is_contiguous(src)) continue;
        if (nelements(t) != nelements(cg->nodes[rk])) continue;
        if (ck < 0 && src == cg->nodes[rk]) { ck = j; continue; }
        if (ck >= 0 && cv < 0 && t->type == cg->nodes[ck]->type && gi.pos(src) < j) {
            const ggml_tensor *v0 = src;
            while (v0 && is_view_op(v0->op) && v0->op != GGML_OP_NONE && v0->src[0]) v0 = v0->src[0];
            if (!v0 || v0->op == GGML_OP_ROPE || is_cpy_op(v0->op)) return false;
            cv = j;
            break;
        }
    }
    if (rk < 0 || ck < 0 || cv < 0) return false;
    out[0] = i; out[1] = rk; out[2] = ck; out[3] = cv;
    return true;
}

static const ggml_tensor *through_views(const ggml_tensor *t) {
    while (t && is_view_op(t->op) && t->op != GGML_OP_NONE && t->src[0]) t = t->src[0];
    return t;
}
// SOFT_MAX at position i followed (views aside) by ARGSORT(desc), GET_ROWS(probs, top-k view), SUM_ROWS, DIV: the MoE router chain of
// llm_build_moe_ffn.  out = positions of the five compute nodes.  The argsort row is only written up to n_used entries by the fused
// kernel, so every reader of the ARGSORT node must be a view of exactly those entries.
static bool match_moe_router(const GraphInfo &gi, int i, const std::vector<char> &done, int out[5]) {
    const ggml_cgraph *cg = gi.cg;
    const ggml_tensor *sm = cg->nodes[i];
    if (sm->op != GGML_OP_SOFT_MAX || sm->src[1] || sm->type != GGML_TYPE_F32 || !is_contiguous(sm) || !sm->src[0] || sm->src[0]->type != GGML_TYPE_F32) return false;
    if (sm->src[0]->nb[0] != 4 || sm->ne[2] != 1 || sm->ne[3] != 1 || sm->ne[0] > 256 || !same_shape(sm, sm->src[0])) return false;
    float scale = 0.f, max_bias = 1.f;
    memcpy(&scale, sm->op_params, 4);
    memcpy(&max_bias, (const float *)sm->op_params + 1, 4);
    if (scale != 1.0f || max_bias != 0.0f) return false;
    // the graph is in depth-first order (ggml_build_forward_expand): the expert matmuls that consume the selection usually sit
    // between ARGSORT and the GET_ROWS / SUM_ROWS / DIV that produce the weights, so search ahead and let can_gather() decide
    int pos[4] = {-1, -1, -1, -1};
    const int horizon = i + 96 < cg->n_nodes ? i + 96 : cg->n_nodes;
    for (int j = i + 1; j < horizon; j++) {
        const ggml_tensor *t = cg->nodes[j];
        if (done[j] || is_view_op(t->op)) continue;
        if (pos[0] < 0 && t->op == GGML_OP_ARGSORT && through_views(t->src[0]) == sm) pos[0] = j;
        else if (pos[0] >= 0 && pos[1] < 0 && t->op == GGML_OP_GET_ROWS && through_views(t->src[0]) == sm && through_views(t->src[1]) == cg->nodes[pos[0]]) pos[1] = j;
        else if (pos[1] >= 0 && pos[2] < 0 && t->op == GGML_OP_SUM_ROWS && through_views(t->src[0]) == cg->nodes[pos[1]]) pos[2] = j;
        else if (pos[2] >= 0 && pos[3] < 0 && t->op == GGML_OP_DIV && through_views(t->src[0]) == cg->nodes[pos[1]] && through_views(t->src[1]) == cg->nodes[pos[2]]) { pos[3] = j; break; }
    }
    if (pos[3] < 0) return false;
    const ggml_tensor *as = cg->nodes[pos[0]], *gr = cg->nodes[pos[1]], *sr = cg->nodes[pos[2]], *dv = cg->nodes[pos[3]];
    const int64_t n_expert = sm->ne[0], n_tok = sm->ne[1];
    if (through_views(as->src[0]) != sm || as->op_params[0] != 1 /* GGML_SORT_ORDER_DESC */ || !is_contiguous(as) || as->ne[0] != n_expert || as->ne[1] != n_tok) return false;
    const ggml_tensor *sel = gr->src[1], *pv = gr->src[0];
    if (!sel || !pv || through_views(pv) != sm || through_views(sel) != as || sel->data != as->data || sel->nb[1] != as->nb[1] || sel->ne[1] != n_tok) return false;
    const int64_t n_used = sel->ne[0];
    if (n_used < 1 || n_used > 32 || n_used > n_expert || pv->ne[0] != 1 || pv->ne[1] != n_expert || pv->data != sm->data) return false;
    if (gr->type != GGML_TYPE_F32 || !is_contiguous(gr) || nelements(gr) != n_used * n_tok) return false;
    if (through_views(sr->src[0]) != gr || sr->src[0]->data != gr->data || !is_contiguous(sr) || nelements(sr) != n_tok || sr->src[0]->ne[0] != n_used) return false;
    if (through_views(dv->src[0]) != gr || dv->src[0]->data != gr->data || through_views(dv->src[1]) != sr || !is_contiguous(dv) || dv->type != GGML_TYPE_F32) return false;
    if (dv->ne[0] != n_used || nelements(dv) != n_used * n_tok) return false;
    // readers of the argsort result: views of the first n_used entries of every row only (followed through further views)
    for (int j = 0; j < cg->n_nodes; j++) {
        const ggml_tensor *t = cg->nodes[j];
        for (int q = 0; q < GGML_MAX_SRC; q++) {
            const ggml_tensor *sq = t->src[q];
            if (!sq || through_views(sq) != as) continue;
            if (is_view_op(t->op)) continue;  // a view of it: checked where it is consumed
            if (sq->data != as->data || sq->ne[0] != n_used || sq->nb[1] != as->nb[1] || sq->ne[1] != n_tok || sq->ne[2] != 1) return false;
        }
    }
    out[0] = i; out[1] = pos[0]; out[2] = pos[1]; out[3] = pos[2]; out[4] = pos[3];
    return true;
}

// Lower cgraph -> Steps.  `done[i]` marks nodes absorbed by a fused step.
static void plan_steps(GraphExec *g, const ggml_cgraph *cg) {
    std::vector<Step> &steps = g->steps;
    steps.clear();
    const int n = cg->n_nodes;
    std::vector<char> done((size_t)n, 0);
    std::vector<int> slot_of((size_t)n, -1);
    int n_slots = 0;
    for (int i = 0; i < n; i++) if (is_cpy_op(cg->nodes[i]->op) && nelements(cg->nodes[i]) != 0) slot_of[(size_t)i] = n_slots++;
    if (!g_use_fusion) {
        for (int i = 0; i < n; i++) {
            const ggml_tensor *a = cg->nodes[i];
            if (is_view_op(a->op) || nelements(a) == 0) continue;
            Step s;
            s.kind = STEP_NODE; s.n = a; s.slot = slot_of[(size_t)i];
            steps.push_back(s);
        }
        return;
    }
    const GraphInfo gi(cg);
    // fused attributes discovered ahead of the matmuls they apply to
    std::unordered_map<const ggml_tensor *, const ggml_tensor *> norm_mul_of;  // activation tensor (MUL node) -> its RMS_NORM node
    std::unordered_map<const ggml_tensor *, const ggml_tensor *> add_of;       // MUL_MAT node -> the ADD node folded into it
    std::unordered_map<int, Step> deferred;                                    // position -> step to emit there
    std::vector<int> exec_pos((size_t)n);                                      // where node i runs: i, or the position of its deferred step
    for (int i = 0; i < n; i++) exec_pos[(size_t)i] = i;
    std::unordered_set<const ggml_tensor *> norm_stored;                       // folded activations whose values a GEMV step already stores

    for (int i = 0; i < n; i++) {
        auto df = deferred.find(i);
        if (df != deferred.end()) { steps.push_back(df->second); continue; }
        if (done[i]) continue;
        const ggml_tensor *a = cg->nodes[i];
        if (is_view_op(a->op) || nelements(a) == 0) continue;
        Step s;
        s.kind = STEP_NODE; s.n = a; s.slot = slot_of[(size_t)i];
        if (is_cpy_op(a->op)) { steps.push_back(s); continue; }

        // the partner of a pair is the next node that is neither a view nor already absorbed
        int j = i + 1;
        while (j < n && (done[j] || is_view_op(cg->nodes[j]->op))) j++;

        if (j < n && rms_mul_pair(a, cg->nodes[j])) {
            const ggml_tensor *b = cg->nodes[j];
            // R1: every reader of the normalised activation is a decode GEMV -> the norm moves into their prologue
            bool all_gemv = gi.sole_user(a) == b && gi.uses(b) > 0 && !(b->flags & GGML_TENSOR_FLAG_OUTPUT) && strncmp(b->name, "result", 6) != 0 &&
                            is_contiguous(a->src[0]) && (((uintptr_t)a->src[0]->data) & 15) == 0;
            const ggml_tensor *w = b->src[0] == a ? b->src[1] : b->src[0];
            all_gemv = all_gemv && (((uintptr_t)w->data) & 15) == 0;
            int readers = 0;
            for (int k = j + 1; k < n && all_gemv; k++) {
                const ggml_tensor *t = cg->nodes[k];
                // the un-normalised input must stay intact until the last GEMV that re-derives the norm from it has run
                if (!is_view_op(t->op) && t->data && ranges_overlap(t->data, nbytes(t), a->src[0]->data, nbytes(a->src[0]))) all_gemv = false;
                for (int q = 0; q < GGML_MAX_SRC; q++) {
                    if (t->src[q] != b) continue;
                    if (q == 1 && mul_mat_fusable(t)) readers++;
                    else all_gemv = false;
                }
                if (readers == gi.uses(b)) break;
            }
            // the folded nodes' values are still stored (by the first GEMV): they must not land on the bytes the GEMVs read
            if (ranges_overlap(a->data, nbytes(a), a->src[0]->data, nbytes(a->src[0])) || ranges_overlap(b->data, nbytes(b), a->src[0]->data, nbytes(a->src[0])) ||
                (((uintptr_t)a->data | (uintptr_t)b->data) & 15) || !is_contiguous(b)) all_gemv = false;
            if (all_gemv && readers == gi.uses(b)) {
                norm_mul_of[b] = a;
                done[j] = 1;
                continue;  // no step: RMS_NORM and MUL are computed inside the GEMVs
            }
            s.kind = STEP_RMS_MUL; s.n2 = b; done[j] = 1; steps.push_back(s);
            continue;
        }
        if (j < n && silu_mul_pair(a, cg->nodes[j])) { s.kind = STEP_SILU_MUL; s.n2 = cg->nodes[j]; done[j] = 1; steps.push_back(s); continue; }
        // prompt graphs order the parallel FFN as gate, SILU, up, MUL (llm_build_ffn through the depth-first graph build): the SILU
        // waits for its MUL two or three nodes further on and the pair runs there as one launch
        if (a->op == GGML_OP_UNARY && a->op_params[0] == GGML_UNARY_OP_SILU && gi.sole_user(a)) {
            const ggml_tensor *mu = gi.sole_user(a);
            const int pm = gi.pos(mu);
            if (pm > i && pm <= i + 6 && !done[pm] && deferred.find(pm) == deferred.end() && silu_mul_pair(a, mu) &&
                can_gather(gi, std::vector<int>{i, pm}, pm, done, exec_pos)) {
                s.kind = STEP_SILU_MUL; s.n2 = mu;
                done[i] = 1; done[pm] = 1;
                exec_pos[(size_t)i] = pm;
                deferred[pm] = s;
                continue;
            }
        }

        // R5: the MoE router chain -> one launch at the position of its SOFT_MAX (the weight nodes move up past the expert matmuls)
        if (a->op == GGML_OP_SOFT_MAX) {
            int m5[5];
            if (match_moe_router(gi, i, done, m5) && can_gather(gi, std::vector<int>(m5, m5 + 5), i, done, exec_pos)) {
                s.kind = STEP_MOE_ROUTER;
                s.group = {cg->nodes[m5[1]], cg->nodes[m5[2]], cg->nodes[m5[3]], cg->nodes[m5[4]]};
                for (int q = 1; q < 5; q++) done[m5[q]] = 1;
                steps.push_back(s);
                continue;
            }
        }

        // R3: a decode-size all-reduce whose only reader is ADD(sum, r) with r already computed -> the residual is added by the
        // exchange kernel itself (peer window, comm.cu); the ADD node disappears
        if (peer_all_reduce_ok(a)) {
            const ggml_tensor *d = gi.sole_user(a);
            if (d && d->op == GGML_OP_ADD && d->type == GGML_TYPE_F32) {
                const ggml_tensor *r = d->src[0] == a ? d->src[1] : d->src[0];
                const int pr = r ? gi.pos(r) : n, pd = gi.pos(d);
                if (r && r != a && r->type == GGML_TYPE_F32 && same_shape(r, a) && same_shape(d, a) && is_contiguous(r) && is_contiguous(d) &&
                    (((uintptr_t)r->data | (uintptr_t)d->data) & 15) == 0 && pr < i && pd > i && !done[pd]) {
                    std::vector<int> mem = {i, pd};
                    if (can_gather(gi, mem, i, done, exec_pos)) {
                        s.kind = STEP_AR_ADD; s.n2 = d; done[pd] = 1;
                        steps.push_back(s);
                        continue;
                    }
                }
            }
        }

        if (a->op == GGML_OP_ROPE) {
            int m[4];
            if (match_rope_kv(gi, i, done, m)) {
                std::vector<int> members(m, m + 4);
                if (can_gather(gi, members, m[3], done, exec_pos)) {
                    Step f;
                    f.kind = STEP_ROPE_KV; f.n = a; f.n2 = cg->nodes[m[1]];
                    f.group = {cg->nodes[m[2]], cg->nodes[m[3]]};
                    f.slot = slot_of[(size_t)m[2]]; f.slot2 = slot_of[(size_t)m[3]];
                    for (int q = 0; q < 4; q++) { done[m[q]] = 1; exec_pos[(size_t)m[q]] = m[3]; }
                    deferred[m[3]] = f;
                    continue;
                }
            }
        }

        if (mul_mat_fusable(a)) {
            // siblings: later MUL_MATs with the same activation tensor, weight type and row stride, hoisted to this position
            s.group.push_back(a);
            const int horizon = i + 48 < n ? i + 48 : n;
            for (int j2 = i + 1; j2 < horizon && s.group.size() < B200_GEMV_MAX_JOBS; j2++) {
                const ggml_tensor *b = cg->nodes[j2];
                if (done[j2] || !mul_mat_fusable(b)) continue;
                if (b->src[1] != a->src[1] || (a->ne[1] > 1 && b->nb[1] != a->nb[1])) continue;
                if (b->src[0]->type != a->src[0]->type) {
                    // decode: a sibling of a second weight type shares the launch (attn_q / attn_k Q4_K + attn_v Q6_K of the *_K_M
                    // mixes), b200_gemv_job.type2: contiguous rows, one of the supported pairs, at most two types per launch
                    if (!g_mixed_types || a->ne[1] != 1 || b->src[0]->ne[0] != a->src[0]->ne[0]) continue;
                    auto kq = [](const ggml_tensor *w) { return w->type == GGML_TYPE_Q4_K || w->type == GGML_TYPE_Q5_K || w->type == GGML_TYPE_Q6_K; };
                    if (!kq(a->src[0]) || !kq(b->src[0])) continue;
                    bool ok = rows_are_contiguous(b->src[0]);
                    for (const ggml_tensor *mm : s.group) {
                        ok = ok && rows_are_contiguous(mm->src[0]);
                        ok = ok && (mm->src[0]->type == a->src[0]->type || mm->src[0]->type == b->src[0]->type);
                    }
                    if (!ok) continue;
                } else if (b->src[0]->nb[1] != a->src[0]->nb[1]) continue;
                std::vector<int> mem = {i, j2};
                if (!can_gather(gi, mem, i, done, exec_pos)) continue;
                s.group.push_back(b);
                done[j2] = 1;
            }
            // R2: a GEMV whose only reader is ADD(mm, r) with r already computed -> residual epilogue, the ADD node disappears
            for (const ggml_tensor *mm : s.group) {
                const ggml_tensor *d = gi.sole_user(mm);
                if (!d || d->op != GGML_OP_ADD || d->type != GGML_TYPE_F32) continue;
                const ggml_tensor *r = d->src[0] == mm ? d->src[1] : d->src[0];
                if (!r || r == mm || r->type != GGML_TYPE_F32 || !same_shape(r, mm) || !same_shape(d, mm)) continue;
                if (!is_contiguous(r) || !is_contiguous(d) || !is_contiguous(mm)) continue;
                const int pr = gi.pos(r), pd = gi.pos(d), pm = gi.pos(mm);
                if (pr >= i || pd < 0 || done[pd]) continue;  // r must exist before the GEMV runs (leafs have pos -1)
                std::vector<int> mem = {pm, pd};
                if (pm != i) mem = {pd};  // a hoisted sibling: only the ADD moves (up to i)
                if (!can_gather(gi, mem, i, done, exec_pos)) continue;
                add_of[mm] = d;
                done[pd] = 1;
            }
            // R4: decode gate/up pair whose results only feed MUL(SILU(gate), up) -> the activation is computed in the epilogue,
            // the SILU and MUL nodes disappear and the two f32 vectors are never written
            if (s.group.size() == 2 && a->ne[1] == 1 && add_of.find(s.group[0]) == add_of.end() && add_of.find(s.group[1]) == add_of.end() &&
                s.group[0]->src[0]->ne[1] == s.group[1]->src[0]->ne[1] && s.group[0]->src[0]->type == s.group[1]->src[0]->type) {
                for (int gi_ = 0; gi_ < 2 && !s.silu_mul; gi_++) {
                    const ggml_tensor *gt = s.group[gi_], *up = s.group[1 - gi_];
                    const ggml_tensor *sn = gi.sole_user(gt);
                    if (!sn || sn->op != GGML_OP_UNARY || sn->op_params[0] != GGML_UNARY_OP_SILU || sn->src[0] != gt) continue;
                    const ggml_tensor *mn = gi.sole_user(sn);
                    if (!mn || mn->op != GGML_OP_MUL || gi.sole_user(up) != mn) continue;
                    if (!((mn->src[0] == sn && mn->src[1] == up) || (mn->src[1] == sn && mn->src[0] == up))) continue;
                    if (!same_shape(mn, gt) || !same_shape(up, gt) || !is_contiguous(mn) || !is_contiguous(gt) || mn->type != GGML_TYPE_F32) continue;
                    const int ps = gi.pos(sn), pm = gi.pos(mn);
                    if (ps < 0 || pm < 0 || done[ps] || done[pm]) continue;
                    // the two matmuls are part of the fused kernel: ggml-alloc usually computes SILU in place over the gate result
                    std::vector<int> mem = {gi.pos(gt), gi.pos(up), ps, pm};
                    std::sort(mem.begin(), mem.end());
                    if (!can_gather(gi, mem, i, done, exec_pos)) continue;
                    s.silu_mul = mn;
                    s.gate_idx = gi_;
                    done[ps] = 1;
                    done[pm] = 1;
                }
            }
            s.kind = STEP_GEMV;
            auto nm = norm_mul_of.find(a->src[1]);
            s.n2 = nm == norm_mul_of.end() ? nullptr : nm->second;  // RMS_NORM node folded into the prologue
            if (s.n2 && norm_stored.insert(a->src[1]).second) {
                // first reader: it also stores the MUL value, and the RMS_NORM value unless its bytes were legitimately reused
                // between the MUL node and this position (it has no reader left then) or alias the MUL output (in-place)
                s.store_norm_mul = true;
                s.store_norm = s.n2->data != a->src[1]->data;
                for (int k2 = gi.pos(a->src[1]) + 1; k2 < i && s.store_norm; k2++) {
                    const ggml_tensor *t = cg->nodes[k2];
                    if (!is_view_op(t->op) && t->data && ranges_overlap(t->data, nbytes(t), s.n2->data, nbytes(s.n2))) s.store_norm = false;
                }
                if (ranges_overlap(s.n2->data, nbytes(s.n2), a->src[1]->data, nbytes(a->src[1]))) s.store_norm = false;
            }
            for (const ggml_tensor *mm : s.group) {
                auto ad = add_of.find(mm);
                s.adds.push_back(ad == add_of.end() ? nullptr : ad->second);
            }
            steps.push_back(s);
            continue;
        }
        steps.push_back(s);
    }
}

// ------------------------------------------------------------------------------------------------ graph key
// Two independent 64-bit hashes over the same words: the first is the cache key, the second is stored with the entry and compared
// on every hit, so that replaying another graph's kernels and pointers needs a simultaneous collision of both (2^-128).
struct KeyHash {
    uint64_t a, b;
};
static inline void mix2(KeyHash &h, uint64_t va, uint64_t vb) {
    h.a ^= va + 0x9e3779b97f4a7c15ull + (h.a << 6) + (h.a >> 2);
    h.a *= 0xff51afd7ed558ccdull;
    h.a ^= h.a >> 32;
    h.b = (h.b ^ vb) * 0x100000001b3ull;  // FNV-1a step on 64-bit words ...
    h.b ^= h.b >> 29;                      // ... with an extra fold (word-sized inputs)
}
static inline void mix(KeyHash &h, uint64_t v) { mix2(h, v, v); }
// The words of ONE node (its descriptor, op_params and source descriptors: at most 118) are first folded into a multilinear digest
// sum_i w_i K_i (mod 2^64) with a fixed odd random K_i per position, once per hash with independent constants: the multiplies do not
// depend on each other, so a 900-node graph hashes in ~35 us instead of ~130 us of serial mixing (B200_HOST_TIMING).  A change of
// any single word changes the digest (K_i is odd); the digest then enters the serial state like any other word.
struct NodeDigest {
    static constexpr int MAXW = 128;
    uint64_t a = 0, b = 0;
    int n = 0;
    static const uint64_t *consts() {
        static uint64_t k[2 * MAXW];
        static const bool init = [] {
            uint64_t x = 0x6a09e667f3bcc908ull;
            for (int i = 0; i < 2 * MAXW; i++) {  // splitmix64
                x += 0x9e3779b97f4a7c15ull;
                uint64_t z = x;
                z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
                z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
                k[i] = (z ^ (z >> 31)) | 1ull;
            }
            return true;
        }();
        (void)init;
        return k;
    }
    inline void add(const uint64_t *k, uint64_t v) {
        const int i = n < MAXW ? n : MAXW - 1;
        a += (v + 0x9e3779b97f4a7c15ull) * k[i];
        b += (v ^ 0xc2b2ae3d27d4eb4full) * k[MAXW + i];
        n++;
    }
};
static inline void hash_tensor(NodeDigest &d, const uint64_t *k, const ggml_tensor *t, bool with_data) {
    d.add(k, (uint64_t)t->type | ((uint64_t)t->op << 32));
    for (int i = 0; i < 4; i++) { d.add(k, (uint64_t)t->ne[i]); d.add(k, (uint64_t)t->nb[i]); }
    if (with_data) d.add(k, (uint64_t)(uintptr_t)t->data);
}
static KeyHash graph_key(const ggml_cgraph *cg) {
    KeyHash h = {0x243f6a8885a308d3ull, 0xcbf29ce484222325ull};
    const uint64_t *k = NodeDigest::consts();
    mix(h, (uint64_t)cg->n_nodes);
    mix(h, (uint64_t)g_use_fusion);
    mix(h, (uint64_t)b200_comm_peer_ready());
    for (int i = 0; i < cg->n_nodes; i++) {
        const ggml_tensor *t = cg->nodes[i];
        if (is_view_op(t->op)) continue;  // views matter only through the tensors that consume them (hashed as srcs)
        const bool cpy = is_cpy_op(t->op);
        NodeDigest d;
        hash_tensor(d, k, t, !cpy);
        const uint64_t *op = (const uint64_t *)t->op_params;
        for (int j = 0; j < GGML_MAX_OP_PARAMS / 8; j++) d.add(k, op[j]);
        for (int j = 0; j < GGML_MAX_SRC; j++) {
            const ggml_tensor *s = t->src[j];
            if (!s) { d.add(k, 0); continue; }
            // CPY(a, b): src[1] is the destination view, its address moves with the KV head and goes through the slot table
            hash_tensor(d, k, s, !(cpy && j == 1));
        }
        mix2(h, d.a ^ (uint64_t)d.n, d.b + (uint64_t)i);
    }
    return h;
}

// ------------------------------------------------------------------------------------------------ execution
static void run_step(GraphExec *g, ExecCtx &x, const Step &s) {
    if (!(s.kind == STEP_NODE && s.n->op == GGML_OP_MUL_MAT)) x.xb_src = nullptr;  // anything else may write the activations or the scratch
    switch (s.kind) {
        case STEP_NODE:
            run_node(x, s.n, s.slot >= 0 ? (void *const *)(g->slots_dev + s.slot) : nullptr);
            return;
        case STEP_MOE_ROUTER: {
            const ggml_tensor *sm = s.n, *as = s.group[0], *gr = s.group[1], *sr = s.group[2], *dv = s.group[3];
            B200_CHECK_RC(b200_op_moe_router((const float *)sm->src[0]->data, (int64_t)(sm->src[0]->nb[1] / 4), (float *)sm->data, (int64_t)(sm->nb[1] / 4),
                                             (int32_t *)as->data, (int64_t)(as->nb[1] / 4), (float *)gr->data, (float *)sr->data, (float *)dv->data,
                                             (int32_t)sm->ne[0], (int32_t)gr->src[1]->ne[0], sm->ne[1], x.stream));
            x.launches++;
            return;
        }
        case STEP_AR_ADD: {
            const ggml_tensor *a = s.n, *d = s.n2;
            const ggml_tensor *r = d->src[0] == a ? d->src[1] : d->src[0];
            B200_CHECK_RC(b200_all_reduce_sum_add_f32((const float *)a->src[0]->data, (const float *)r->data, (float *)d->data, nelements(d), x.stream));
            x.launches++;
            return;
        }
        case STEP_RMS_MUL: {
            const ggml_tensor *a = s.n, *b = s.n2;
            const ggml_tensor *w = b->src[0] == a ? b->src[1] : b->src[0];
            float eps;
            memcpy(&eps, a->op_params, 4);
            b200_tensor tx = as_b200(a->src[0]), td = as_b200(b);
            // the RMS_NORM node's own value is still written: another node may read it, and ggml-alloc may have aliased it
            // with the MUL output (in-place), in which case norm_out == dst and the kernel writes dst last
            float *norm_out = a->data == b->data ? nullptr : (float *)a->data;
            B200_CHECK_RC(b200_op_rms_norm_mul(&tx, (const float *)w->data, norm_out, &td, eps, x.stream));
            x.launches++;
            return;
        }
        case STEP_SILU_MUL: {
            const ggml_tensor *a = s.n, *b = s.n2;
            const ggml_tensor *u = b->src[0] == a ? b->src[1] : b->src[0];
            float *act_out = a->data == b->data ? nullptr : (float *)a->data;
            B200_CHECK_RC(b200_op_silu_mul((const float *)a->src[0]->data, (const float *)u->data, act_out, (float *)b->data, nelements(b), x.stream));
            x.launches++;
            return;
        }
        case STEP_GEMV: {
            b200_gemv_job jobs[B200_GEMV_MAX_JOBS];
            b200_gemv_fuse fuse;
            memset(&fuse, 0, sizeof(fuse));
            const ggml_tensor *a = s.group[0];
            const ggml_tensor *act = a->src[1];
            const float *X = (const float *)act->data;
            int64_t x_stride = (int64_t)(act->nb[1] / 4);
            if (s.n2) {  // RMS_NORM -> MUL folded into the prologue: read the un-normalised rows
                const ggml_tensor *rn = s.n2, *mul = act;
                const ggml_tensor *w = mul->src[0] == rn ? mul->src[1] : mul->src[0];
                X = (const float *)rn->src[0]->data;
                x_stride = (int64_t)(rn->src[0]->nb[1] / 4);
                fuse.norm_weight = (const float *)w->data;
                memcpy(&fuse.norm_eps, rn->op_params, 4);
            }
            int nj = 0;
            if (s.silu_mul) {  // jobs[0] = gate, jobs[1] = up, one output vector
                for (int q = 0; q < 2; q++) {
                    const ggml_tensor *m = s.group[q == 0 ? s.gate_idx : 1 - s.gate_idx];
                    jobs[q].W = m->src[0]->data; jobs[q].X = X; jobs[q].M = m->src[0]->ne[1]; jobs[q].id_index = -1; jobs[q]._pad = 0;
                    jobs[q].dst = (float *)s.silu_mul->data;
                }
                fuse.silu_pair = 1;
                {   // the matmul / SILU nodes keep their values unless ggml-alloc placed a later node of the group on the same bytes
                    const ggml_tensor *gt = s.group[s.gate_idx], *up = s.group[1 - s.gate_idx], *mn = s.silu_mul;
                    const ggml_tensor *sn = mn->src[0] == up ? mn->src[1] : mn->src[0];
                    auto clash = [&](const ggml_tensor *t, std::initializer_list<const ggml_tensor *> later) {
                        for (const ggml_tensor *o : later) if (ranges_overlap(t->data, nbytes(t), o->data, nbytes(o))) return true;
                        return false;
                    };
                    fuse.gate_out = clash(gt, {up, sn, mn}) ? nullptr : (float *)gt->data;
                    fuse.up_out = clash(up, {sn, mn}) ? nullptr : (float *)up->data;
                    fuse.silu_out = clash(sn, {mn}) ? nullptr : (float *)sn->data;
                }
                if (s.n2) {
                    fuse.norm_out = s.store_norm ? (float *)s.n2->data : nullptr;
                    fuse.norm_mul_out = s.store_norm_mul ? (float *)act->data : nullptr;
                }
                B200_CHECK_RC(b200_mul_mat_vec_q_fused(a->src[0]->type, jobs, 2, a->src[0]->nb[1], x_stride, (int64_t)(a->nb[1] / 4), a->src[0]->ne[0], 1, &fuse, x.stream));
                x.launches++;
                return;
            }
            for (size_t q = 0; q < s.group.size(); q++) {
                const ggml_tensor *m = s.group[q], *add = s.adds[q];
                jobs[nj].W = m->src[0]->data; jobs[nj].X = X; jobs[nj].M = m->src[0]->ne[1]; jobs[nj].id_index = -1;
                jobs[nj]._pad = m->src[0]->type != a->src[0]->type ? (int32_t)m->src[0]->type : 0;  // a member of the second weight type
                if (add) {
                    const ggml_tensor *r = add->src[0] == m ? add->src[1] : add->src[0];
                    jobs[nj].dst = (float *)add->data;
                    fuse.residual[nj] = (const float *)r->data;
                    fuse.raw_dst[nj] = ranges_overlap(m->data, nbytes(m), add->data, nbytes(add)) ? nullptr : (float *)m->data;
                } else {
                    jobs[nj].dst = (float *)m->data;
                }
                nj++;
            }
            if (s.n2) {
                fuse.norm_out = s.store_norm ? (float *)s.n2->data : nullptr;
                fuse.norm_mul_out = s.store_norm_mul ? (float *)act->data : nullptr;
            }
            B200_CHECK_RC(b200_mul_mat_vec_q_fused(a->src[0]->type, jobs, nj, a->src[0]->nb[1], x_stride, (int64_t)(a->nb[1] / 4), a->src[0]->ne[0], act->ne[1],
                                                   &fuse, x.stream));
            x.launches++;
            return;
        }
        case STEP_ROPE_KV: {
            const ggml_tensor *rq = s.n, *rk = s.n2, *ck = s.group[0], *cv = s.group[1];
            b200_rope_params rp;
            rp.n_dims = rq->op_params[1]; rp.mode = rq->op_params[2]; rp.n_ctx_orig = rq->op_params[4];
            memcpy(&rp.freq_base, rq->op_params + 5, 4); memcpy(&rp.freq_scale, rq->op_params + 6, 4); memcpy(&rp.ext_factor, rq->op_params + 7, 4);
            memcpy(&rp.attn_factor, rq->op_params + 8, 4); memcpy(&rp.beta_fast, rq->op_params + 9, 4); memcpy(&rp.beta_slow, rq->op_params + 10, 4);
            b200_tensor q = as_b200(rq->src[0]), qd = as_b200(rq), k = as_b200(rk->src[0]), kd = as_b200(rk), v = as_b200(cv->src[0]);
            B200_CHECK_RC(b200_op_rope_kv_store(&q, &qd, &k, &kd, &v, (const int32_t *)rq->src[1]->data, rq->src[2] ? (const float *)rq->src[2]->data : nullptr, &rp,
                                                ck->type, nullptr, nullptr, (void *const *)(g->slots_dev + s.slot), (void *const *)(g->slots_dev + s.slot2), x.stream));
            x.launches++;
            return;
        }
    }
}

static void ensure_workspaces(GraphExec *g, const ggml_cgraph *cg, int n_slots) {
    size_t fa = 0, sc = 0;
    for (int i = 0; i < cg->n_nodes; i++) {
        const ggml_tensor *t = cg->nodes[i];
        if (t->op == GGML_OP_FLASH_ATTN_EXT) { const size_t b = node_fa_workspace(t); if (b > fa) fa = b; }
        if (t->op == GGML_OP_MUL_MAT || t->op == GGML_OP_MUL_MAT_ID) { const size_t b = node_scratch_bytes(t); if (b > sc) sc = b; }
    }
    // growing a workspace moves it: every captured graph has the old pointer baked in
    if (fa > g->fa_ws_bytes) {
        B200_CHECK_CUDA(cudaStreamSynchronize(g->stream));
        drop_cache(g);
        if (g->fa_ws) B200_CHECK_CUDA(cudaFree(g->fa_ws));
        fa += fa / 4;
        B200_CHECK_CUDA(cudaMalloc(&g->fa_ws, fa));
        B200_CHECK_CUDA(cudaMemsetAsync(g->fa_ws, 0, fa, g->stream));
        g->fa_ws_bytes = fa;
    }
    if (sc > g->scratch_bytes) {
        B200_CHECK_CUDA(cudaStreamSynchronize(g->stream));
        drop_cache(g);
        if (g->scratch) B200_CHECK_CUDA(cudaFree(g->scratch));
        B200_CHECK_CUDA(cudaMalloc(&g->scratch, sc));
        g->scratch_bytes = sc;
    }
    if (n_slots > g->slots_cap) {
        B200_CHECK_CUDA(cudaStreamSynchronize(g->stream));
        drop_cache(g);
        if (g->slots_dev) B200_CHECK_CUDA(cudaFree(g->slots_dev));
        g->slots_cap = n_slots + 64;
        B200_CHECK_CUDA(cudaMalloc((void **)&g->slots_dev, sizeof(void *) * (size_t)g->slots_cap));
    }
}

static int count_cpy(const ggml_cgraph *cg) {
    int c = 0;
    for (int i = 0; i < cg->n_nodes; i++) if (is_cpy_op(cg->nodes[i]->op) && nelements(cg->nodes[i]) != 0) c++;
    return c;
}

// refresh the CPY destination slots (pageable source: the runtime stages it before returning, so the vector is reusable)
static void upload_slots(GraphExec *g, const ggml_cgraph *cg, int n_slots) {
    if (n_slots == 0) return;
    g->slots_host.resize((size_t)n_slots);
    int k = 0;
    for (int i = 0; i < cg->n_nodes; i++) {
        const ggml_tensor *t = cg->nodes[i];
        if (is_cpy_op(t->op) && nelements(t) != 0) g->slots_host[(size_t)k++] = t->data;
    }
    B200_CHECK_CUDA(cudaMemcpyAsync(g->slots_dev, g->slots_host.data(), sizeof(void *) * (size_t)n_slots, cudaMemcpyHostToDevice, g->stream));
}

// B200_DRY_RUN: plan the graph, account the launches, check that every node is one supports_op accepted, zero the outputs
static const char *step_kind_name(int k) {
    static const char *names[] = {"NODE", "RMS_MUL", "SILU_MUL", "GEMV", "ROPE_KV", "AR_ADD", "MOE_ROUTER"};
    return k >= 0 && k < (int)(sizeof(names) / sizeof(names[0])) ? names[k] : "?";
}
static void dry_compute(GraphExec *g, const ggml_cgraph *cg) {
    plan_steps(g, cg);
    static const bool dump = getenv("B200_PLAN_DUMP") != nullptr;
    for (int i = 0; i < cg->n_nodes; i++) {
        const ggml_tensor *t = cg->nodes[i];
        if (!op_supported(g->device, t)) {
            fprintf(stderr, "B200 dry run: node %d (%s, op %d) reached graph_compute but supports_op rejects it\n", i, t->name, (int)t->op);
            backend_abort("graph_compute", "unsupported node");
        }
        if (!is_view_op(t->op) && t->data && !(t->buffer && buft_is_b200_split(t->buffer->buft))) memset(t->data, 0, nbytes(t));
    }
    int64_t launches = 0;
    for (const Step &s : g->steps) {
        int64_t l = 1;
        if (s.kind == STEP_NODE && s.n->op == GGML_OP_MUL_MAT && s.n->src[0] && is_quantized(s.n->src[0]->type) && s.n->src[1]->ne[1] > 8) l = 1;
        launches += l;
        if (dump) {
            fprintf(stderr, "  step %-10s %-24s", step_kind_name(s.kind), s.n ? s.n->name : "");
            for (const ggml_tensor *m : s.group) fprintf(stderr, " +%s", m->name);
            if (s.n2) fprintf(stderr, " n2=%s", s.n2->name);
            if (s.silu_mul) fprintf(stderr, " silu_mul=%s", s.silu_mul->name);
            for (const ggml_tensor *ad : s.adds) if (ad) fprintf(stderr, " add=%s", ad->name);
            if (s.store_norm_mul) fprintf(stderr, " stores_norm%s", s.store_norm ? "+rms" : "");
            fprintf(stderr, "\n");
        }
    }
    if (dump) fprintf(stderr, "B200 dry run: graph of %d nodes -> %zu steps\n", cg->n_nodes, g->steps.size());
    g_stat_launches += launches;
    g_stat_replays++;
}

// B200_HOST_TIMING=1: host time spent per graph_compute call before the launch (slot upload, key hash), printed at exit.  In a dry
// run (no GPU) this separates the backend's per-token host cost from the caller's own graph building and scheduling.
static const bool g_host_timing = getenv("B200_HOST_TIMING") != nullptr;
static double g_ht_prelude = 0, g_ht_key = 0, g_ht_dry = 0;
static long g_ht_calls = 0;
static double ht_now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static void ht_report() {
    fprintf(stderr, "B200 host timing: %ld graph_compute calls; per call: prelude %.1f us, key %.1f us, dry-run planning %.1f us\n", g_ht_calls,
            1e6 * g_ht_prelude / (double)(g_ht_calls ? g_ht_calls : 1), 1e6 * g_ht_key / (double)(g_ht_calls ? g_ht_calls : 1),
            1e6 * g_ht_dry / (double)(g_ht_calls ? g_ht_calls : 1));
}

enum ggml_status graph_exec_compute(GraphExec *g, ggml_cgraph *cg) {
    if (cg->n_nodes == 0) return GGML_STATUS_SUCCESS;
    const double ht0 = g_host_timing ? ht_now() : 0;
    if (g_host_timing && g_ht_calls++ == 0) atexit(ht_report);
    const int n_slots = count_cpy(cg);
    ensure_workspaces(g, cg, n_slots);
    upload_slots(g, cg, n_slots);
    ExecCtx x;
    x.device = g->device; x.stream = g->stream; x.scratch = g->scratch; x.scratch_bytes = g->scratch_bytes; x.fa_ws = g->fa_ws; x.fa_ws_bytes = g->fa_ws_bytes; x.launches = 0;

    // graphs with a handful of nodes (op tests, K-shift) are not worth an instantiate
    // row-split weights fan a matmul out over several devices and streams: run those graphs eagerly (the reference also
    // turns CUDA graphs off for split buffers, ggml-cuda.cu:2561-2568)
    bool has_split = false;
    for (int i = 0; i < cg->n_nodes && !has_split; i++) {
        const ggml_tensor *w = cg->nodes[i]->src[0];
        has_split = cg->nodes[i]->op == GGML_OP_MUL_MAT && w && w->buffer && buft_is_b200_split(w->buffer->buft);
    }
    if (g_host_timing) {
        const double t1 = ht_now();
        g_ht_prelude += t1 - ht0;
        if (g_dry) { (void)graph_key(cg); g_ht_key += ht_now() - t1; }
    }
    if (g_dry) {
        const double t2 = g_host_timing ? ht_now() : 0;
        dry_compute(g, cg);
        if (g_host_timing) g_ht_dry += ht_now() - t2;
        return GGML_STATUS_SUCCESS;
    }
    const bool use_graph = g_use_graphs && cg->n_nodes >= 16 && !has_split;
    if (!use_graph) {
        plan_steps(g, cg);
        for (const Step &s : g->steps) run_step(g, x, s);
        g_stat_launches += x.launches;
        return GGML_STATUS_SUCCESS;
    }
    const double t3 = g_host_timing ? ht_now() : 0;
    const KeyHash kh = graph_key(cg);
    if (g_host_timing) g_ht_key += ht_now() - t3;
    const uint64_t key = kh.a;
    g->tick++;
    auto it = g->cache.find(key);
    if (it != g->cache.end() && it->second.check != kh.b) {  // same key, different graph: drop the entry and capture afresh
        B200_CHECK_CUDA(cudaStreamSynchronize(g->stream));
        cudaGraphExecDestroy(it->second.exec);
        g->cache.erase(it);
        it = g->cache.end();
    }
    if (it == g->cache.end()) {
        if (g->cache.size() >= 24) {  // prefill with ever-changing batch sizes must not leak executables: evict the oldest
            auto old = g->cache.begin();
            for (auto c = g->cache.begin(); c != g->cache.end(); ++c) if (c->second.last_used < old->second.last_used) old = c;
            B200_CHECK_CUDA(cudaStreamSynchronize(g->stream));
            cudaGraphExecDestroy(old->second.exec);
            g->cache.erase(old);
        }
        plan_steps(g, cg);
        cudaGraph_t graph = nullptr;
        B200_CHECK_CUDA(cudaStreamBeginCapture(g->stream, cudaStreamCaptureModeThreadLocal));
        for (const Step &s : g->steps) run_step(g, x, s);
        B200_CHECK_CUDA(cudaStreamEndCapture(g->stream, &graph));
        CachedGraph c;
        B200_CHECK_CUDA(cudaGraphInstantiate(&c.exec, graph, 0));
        B200_CHECK_CUDA(cudaGraphDestroy(graph));
        c.launches = x.launches;
        c.n_slots = n_slots;
        c.check = kh.b;
        it = g->cache.emplace(key, c).first;
        g_stat_captures++;
    }
    it->second.last_used = g->tick;
    B200_CHECK_CUDA(cudaGraphLaunch(it->second.exec, g->stream));
    g_stat_replays++;
    g_stat_launches += it->second.launches;
    return GGML_STATUS_SUCCESS;
}

}  // namespace b200
