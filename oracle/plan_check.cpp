// oracle/plan_check.cpp — TEST INFRASTRUCTURE.  Runs the unmodified reference (CPU backend) on a GGUF model, rebuilds
// the single-token decode graph it executed (same node order, real ggml-alloc addresses) and feeds it to the plugin's
// chain planner (csrc/backend/chain_plan.cpp).  Prints one line per plan item so tests can check that the whole decode
// graph is recognised.  No GPU needed: the planner is host logic.
#include "llama.h"
#include "ggml.h"
#include "ggml-backend.h"
#include "b200_plugin.h"

#include <cstdio>
#include <vector>

static std::vector<ggml_tensor *> g_nodes;
static bool g_collect = false;
static bool on_node(struct ggml_tensor * t, bool ask, void *) {
    if (ask && g_collect) g_nodes.push_back(t);
    return false;
}

int main(int argc, char ** argv) {
    if (argc < 3) { fprintf(stderr, "usage: plan-check model.gguf backend_dir\n"); return 2; }
    ggml_backend_load_all_from_path(argv[2]);
    llama_model_params mp = llama_model_default_params();
    mp.n_gpu_layers = 0;
    llama_model * m = llama_model_load_from_file(argv[1], mp);
    if (!m) return 3;
    llama_context_params cp = llama_context_default_params();
    cp.n_ctx = 512; cp.n_batch = 512; cp.flash_attn_type = LLAMA_FLASH_ATTN_TYPE_ENABLED;
    cp.cb_eval = on_node; cp.n_threads = 4;
    llama_context * c = llama_init_from_model(m, cp);
    std::vector<llama_token> toks = {1, 2, 3, 4};
    llama_decode(c, llama_batch_get_one(toks.data(), 4));
    llama_token t = 5;
    llama_decode(c, llama_batch_get_one(&t, 1));
    g_collect = true;
    llama_decode(c, llama_batch_get_one(&t, 1));
    g_collect = false;

    ggml_init_params ip = { ggml_graph_overhead_custom(g_nodes.size() + 64, false) + ggml_tensor_overhead() * 8, nullptr, true };
    ggml_context * ctx = ggml_init(ip);
    ggml_cgraph * gf = ggml_new_graph_custom(ctx, g_nodes.size() + 64, false);
    for (ggml_tensor * n : g_nodes) ggml_build_forward_expand(gf, n);       // same order: sources were visited first
    bool same_order = (size_t) ggml_graph_n_nodes(gf) == g_nodes.size();
    for (int i = 0; same_order && i < ggml_graph_n_nodes(gf); ++i) same_order = ggml_graph_node(gf, i) == g_nodes[(size_t) i];
    printf("nodes %d same_order %d\n", ggml_graph_n_nodes(gf), (int) same_order);

    b200_plan plan;
    b200_plan_build(gf, plan);
    int phases = 0, attn = 0, plain = 0;
    for (const b200_plan_item & it : plan.items) {
        const ggml_tensor * n = ggml_graph_node(gf, it.first_node);
        if (it.kind == B200_ITEM_PHASE) {
            ++phases; attn += it.phase.kind == B200_PHASE_ATTN;
            if (it.phase.kind == B200_PHASE_GEMV)
                printf("PHASE gemv node %4d %-24s K=%lld segs=%d norm=%d glu=%d bias=%d\n", it.first_node, n->name, (long long) it.phase.gemv.x.ne[0],
                       it.phase.gemv.n_seg, it.phase.gemv.norm_w != nullptr, it.phase.gemv.glu_op, it.phase.gemv.seg[it.phase.gemv.n_seg - 1].bias != nullptr);
            else
                printf("PHASE attn node %4d %-24s n_kv=%lld\n", it.first_node, n->name, (long long) it.phase.attn.k_cache.ne[1]);
        } else {
            const bool noop = n->op == GGML_OP_VIEW || n->op == GGML_OP_RESHAPE || n->op == GGML_OP_PERMUTE || n->op == GGML_OP_TRANSPOSE || n->op == GGML_OP_NONE;
            if (!noop) { ++plain; printf("NODE  %-12s node %4d %s\n", ggml_op_name(n->op), it.first_node, n->name); }
        }
    }
    printf("summary phases %d attn %d plain_compute_nodes %d scratch_floats %lld\n", phases, attn, plain, (long long) plan.scratch_floats);
    return 0;
}
