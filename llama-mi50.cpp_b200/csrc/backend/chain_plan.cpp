// chain_plan.cpp — compiles the single-token part of a ggml graph into decode-chain phases.
//
// The reference executes a decode graph node by node with local fusions (ggml-cuda.cu:3394-3922: rms_norm+mul,
// gate/up+GLU, rope+set_rows, ...).  Here whole runs of nodes become PHASES of one persistent kernel
// (include/b200_kernels.h, b200_chain_*):
//   A  [RMS_NORM, MUL, MUL_MAT q/k/v (+ADD bias), ROPE q, ROPE k, SET_ROWS k, SET_ROWS v, FLASH_ATTN_EXT]
//        -> GEMV phase (norm fused, 3 row segments, outputs in a private scratch) + attention phase
//   G1 [RMS_NORM, MUL, MUL_MAT x n (+ADD bias) | MUL_MAT gate, MUL_MAT up, GLU]  -> GEMV phase with the norm fused
//   G2 [MUL_MAT (+ADD bias | residual)]                                          -> GEMV phase
// The graph is the llama.cpp decode graph of src/models/llama.cpp:3-165 + llama-graph.cpp:1675-1900 (and every
// architecture built from the same pieces); nothing is assumed about node ORDER beyond what ggml guarantees
// (sources precede users): patterns are recognised by data flow and every elided tensor is checked to have all of
// its uses inside the pattern.  Anything unrecognised stays a plain node and simply ends the current chain.
#include "b200_plugin.h"

#include <algorithm>
#include <cstring>
#include <unordered_map>

namespace {

bool is_view_op(ggml_op op) { return op == GGML_OP_VIEW || op == GGML_OP_RESHAPE || op == GGML_OP_PERMUTE || op == GGML_OP_TRANSPOSE; }

bool is_f32_vec(const ggml_tensor * t, int64_t n) {
    return t && t->type == GGML_TYPE_F32 && t->nb[0] == sizeof(float) && t->ne[0] == n && ggml_nelements(t) == n;
}

// GET_ROWS of the only row of a one-row tensor (llama.cpp's inp_out_ids on the last layer of a 1-token batch)
bool is_identity_get_rows(const ggml_tensor * t) {
    if (t->op != GGML_OP_GET_ROWS) return false;
    const ggml_tensor * a = t->src[0]; const ggml_tensor * ids = t->src[1];
    return a->type == GGML_TYPE_F32 && ggml_nelements(a) == a->ne[0] && ids->type == GGML_TYPE_I32 && ggml_nelements(ids) == 1 &&
           is_f32_vec(t, a->ne[0]) && a->nb[0] == sizeof(float);
}

bool overlap(const void * a, size_t na, const void * b, size_t nb) {
    const char * pa = (const char *) a; const char * pb = (const char *) b;
    return pa < pb + nb && pb < pa + na;
}

b200_tensor vec_desc(void * data, int64_t n) {
    b200_tensor d{};
    d.data = data; d.type = B200_F32;
    d.ne[0] = n; d.ne[1] = d.ne[2] = d.ne[3] = 1;
    d.nb[0] = 4; d.nb[1] = d.nb[2] = d.nb[3] = 4 * n;
    return d;
}

struct matcher {
    ggml_cgraph * g;
    b200_plan & plan;
    std::unordered_map<const ggml_tensor *, int> index;
    std::unordered_map<const ggml_tensor *, void *> redirect;      // tensor -> plan-private (virtual) address
    std::unordered_map<const ggml_tensor *, int> redirect_uses;
    bool allow_redirect = true;

    // descriptor of a tensor a phase READS: follows a redirect and counts the use
    b200_tensor read_desc(const ggml_tensor * t) {
        b200_tensor d = b200_desc(t);
        auto it = redirect.find(t);
        if (it != redirect.end()) d.data = it->second;
        return d;
    }
    void * valloc(int64_t bytes) {
        void * p = (void *) (B200_REDIRECT_VBASE + (uintptr_t) plan.redirect_bytes);
        plan.redirect_bytes += (bytes + 255) & ~(int64_t) 255;
        return p;
    }
    // every redirected tensor must be read by chain phases only: count the reads in the finished plan
    bool redirects_ok() {
        std::unordered_map<const void *, const ggml_tensor *> by_addr;
        for (auto & kv : redirect) by_addr[kv.second] = kv.first;
        redirect_uses.clear();
        auto note = [&](const void * p) { auto it = by_addr.find(p); if (it != by_addr.end()) redirect_uses[it->second]++; };
        for (const b200_plan_item & it : plan.items) {
            if (it.kind != B200_ITEM_PHASE || it.phase.kind != B200_PHASE_GEMV) continue;
            note(it.phase.gemv.x.data);
            for (int s = 0; s < it.phase.gemv.n_seg; ++s) if (it.phase.gemv.seg[s].bias) note(it.phase.gemv.seg[s].bias->data);
        }
        for (auto & kv : redirect) {
            const int i = idx(kv.first);
            auto u = redirect_uses.find(kv.first);
            if (i < 0 || is_output(kv.first) || ggml_node_get_use_count(g, i) != (u == redirect_uses.end() ? 0 : u->second)) return false;
        }
        return true;
    }

    matcher(ggml_cgraph * g_, b200_plan & p) : g(g_), plan(p) {
        for (int i = 0; i < g->n_nodes; ++i) index[g->nodes[i]] = i;
    }
    int idx(const ggml_tensor * t) const { auto it = index.find(t); return it == index.end() ? -1 : it->second; }
    bool is_output(const ggml_tensor * t) const { return (t->flags & GGML_TENSOR_FLAG_OUTPUT) != 0; }
    int uses(const ggml_tensor * t) const { const int i = idx(t); return i < 0 ? -1 : ggml_node_get_use_count(g, i); }

    const b200_tensor * keep(const ggml_tensor * t) { plan.pool.push_back(b200_desc(t)); return &plan.pool.back(); }

    // every node of `elided` (never materialised) must be used only from inside `members`
    bool uses_confined(std::vector<int> members, std::vector<int> elided) const {
        std::sort(members.begin(), members.end()); members.erase(std::unique(members.begin(), members.end()), members.end());
        std::sort(elided.begin(), elided.end());   elided.erase(std::unique(elided.begin(), elided.end()), elided.end());
        for (int e : elided) {
            const ggml_tensor * t = g->nodes[e];
            if (is_output(t)) return false;
            int refs = 0;
            for (int m : members)
                for (int s = 0; s < GGML_MAX_SRC; ++s) refs += g->nodes[m]->src[s] == t;
            if (refs != ggml_node_get_use_count(g, e)) return false;
        }
        return true;
    }

    bool single_column_mm(const ggml_tensor * n) const {
        if (n->op != GGML_OP_MUL_MAT || n->type != GGML_TYPE_F32) return false;
        const ggml_tensor * w = n->src[0]; const ggml_tensor * x = n->src[1];
        return ggml_nelements(n) == n->ne[0] && n->nb[0] == sizeof(float) && w->ne[2] == 1 && w->ne[3] == 1 && n->ne[0] == w->ne[1] &&
               is_f32_vec(x, w->ne[0]) && ggml_is_quantized(w->type);
    }

    // [RMS_NORM, MUL(norm, w)] at i: fills x / norm_w / eps, returns the MUL node or null
    const ggml_tensor * match_norm(int i, b200_gemv_fused_args & a, bool allow_output = false) {
        if (i + 1 >= g->n_nodes) return nullptr;
        const ggml_tensor * rn = g->nodes[i]; const ggml_tensor * mul = g->nodes[i + 1];
        if (rn->op != GGML_OP_RMS_NORM || mul->op != GGML_OP_MUL) return nullptr;
        const ggml_tensor * x = rn->src[0];
        const int64_t K = x->ne[0];
        if (!is_f32_vec(x, K) || !is_f32_vec(rn, K) || !is_f32_vec(mul, K)) return nullptr;
        const ggml_tensor * w = mul->src[0] == rn ? mul->src[1] : mul->src[1] == rn ? mul->src[0] : nullptr;
        if (!w || !is_f32_vec(w, K) || w->op != GGML_OP_NONE) return nullptr;
        if (ggml_node_get_use_count(g, i) != 1 || is_output(rn) || (is_output(mul) && !allow_output)) return nullptr;
        a.x = read_desc(x);
        a.norm_w = keep(w);
        memcpy(&a.eps, rn->op_params, sizeof(float));
        return mul;
    }

    // a MUL_MAT optionally followed through its single use by ADD(bias): returns the tensor carrying the result
    // (the ADD, or the MUL_MAT itself); nodes are appended to `members`, the MUL_MAT to `elided` when an ADD follows
    const ggml_tensor * mm_with_bias(const ggml_tensor * mm, const ggml_tensor *& bias, std::vector<int> & members, std::vector<int> & elided) {
        bias = nullptr;
        const int i = idx(mm);
        members.push_back(i);
        if (ggml_node_get_use_count(g, i) != 1 || is_output(mm)) return mm;
        for (int j = i + 1; j < g->n_nodes && j <= i + 4; ++j) {
            const ggml_tensor * n = g->nodes[j];
            if (n->op != GGML_OP_ADD || (n->src[0] != mm && n->src[1] != mm)) continue;
            const ggml_tensor * other = n->src[0] == mm ? n->src[1] : n->src[0];
            if (!is_f32_vec(other, mm->ne[0]) || !is_f32_vec(n, mm->ne[0]) || other->op != GGML_OP_NONE) return mm;   // only weight biases here
            bias = other;
            members.push_back(j); elided.push_back(i);
            return n;
        }
        return mm;
    }

    // ---------------------------------------------------------------- pattern A: attention block
    int match_attention(int i) {
        b200_phase pq{}; pq.kind = B200_PHASE_GEMV;
        const ggml_tensor * mul = match_norm(i, pq.gemv);
        if (!mul) return 0;
        // the first FLASH_ATTN_EXT within reach
        int f = -1;
        for (int j = i + 2; j < g->n_nodes && j < i + 48; ++j)
            if (g->nodes[j]->op == GGML_OP_FLASH_ATTN_EXT) { f = j; break; }
        if (f < 0) return 0;
        const ggml_tensor * fa = g->nodes[f];
        if (fa->type != GGML_TYPE_F32) return 0;
        const ggml_tensor * sinks = fa->src[4];
        if (sinks && (sinks->type != GGML_TYPE_F32 || sinks->op != GGML_OP_NONE || sinks->nb[0] != sizeof(float))) return 0;
        float scale, max_bias, softcap;
        memcpy(&scale, (const float *) fa->op_params + 0, 4); memcpy(&max_bias, (const float *) fa->op_params + 1, 4); memcpy(&softcap, (const float *) fa->op_params + 2, 4);
        if (max_bias != 0.0f || softcap != 0.0f) return 0;

        std::vector<int> members = {i, i + 1, f}, elided = {i, i + 1};
        auto through_views = [&](const ggml_tensor * t, bool add) -> const ggml_tensor * {
            while (t && is_view_op(t->op)) {
                const int k = idx(t);
                if (k < i || k > f) return nullptr;                      // a view created outside the block: not ours to judge
                if (add) { members.push_back(k); elided.push_back(k); }
                t = t->src[0];
            }
            return t;
        };
        // ---- q: FA.src[0] -> views -> ROPE -> views -> [ADD] -> MUL_MAT
        const ggml_tensor * q = fa->src[0];
        const ggml_tensor * rq = through_views(q, true);
        if (!rq || rq->op != GGML_OP_ROPE || idx(rq) < i) return 0;
        members.push_back(idx(rq)); elided.push_back(idx(rq));
        const ggml_tensor * q_lin = through_views(rq->src[0], true);
        // ---- k, v: cache views of FA -> leaf cache tensors; the SET_ROWS nodes that write those caches
        const ggml_tensor * kc = fa->src[1]; const ggml_tensor * vc = fa->src[2];
        const ggml_tensor * k_leaf = through_views(kc, true); const ggml_tensor * v_leaf = through_views(vc, true);
        if (!q_lin || !k_leaf || !v_leaf || k_leaf->op != GGML_OP_NONE || v_leaf->op != GGML_OP_NONE) return 0;
        const ggml_tensor * sr_k = nullptr; const ggml_tensor * sr_v = nullptr;
        for (int j = i + 2; j < f; ++j) {
            const ggml_tensor * n = g->nodes[j];
            if (n->op != GGML_OP_SET_ROWS) continue;
            const ggml_tensor * target = n->src[2] ? n->src[2] : n->view_src;
            while (target && is_view_op(target->op)) target = target->src[0];
            if (target == k_leaf && !sr_k) sr_k = n; else if (target == v_leaf && !sr_v) sr_v = n; else return 0;
        }
        if (!sr_k || !sr_v) return 0;
        members.push_back(idx(sr_k)); members.push_back(idx(sr_v));
        const ggml_tensor * rk = through_views(sr_k->src[0], true);
        if (!rk || rk->op != GGML_OP_ROPE || idx(rk) < i) return 0;
        members.push_back(idx(rk)); elided.push_back(idx(rk));
        const ggml_tensor * k_lin = through_views(rk->src[0], true);
        const ggml_tensor * v_lin = through_views(sr_v->src[0], true);
        if (!k_lin || !v_lin) return 0;

        // ---- the three projections
        const ggml_tensor * lin[3] = {q_lin, k_lin, v_lin};
        const ggml_tensor * bias[3] = {nullptr, nullptr, nullptr};
        const ggml_tensor * mm[3];
        for (int s = 0; s < 3; ++s) {
            const ggml_tensor * t = lin[s];
            if (t->op == GGML_OP_ADD) {
                const ggml_tensor * a0 = t->src[0]; const ggml_tensor * a1 = t->src[1];
                const ggml_tensor * m = a0->op == GGML_OP_MUL_MAT ? a0 : a1->op == GGML_OP_MUL_MAT ? a1 : nullptr;
                if (!m) return 0;
                bias[s] = m == a0 ? a1 : a0;
                if (!is_f32_vec(bias[s], m->ne[0]) || bias[s]->op != GGML_OP_NONE || idx(t) < i) return 0;
                members.push_back(idx(t)); elided.push_back(idx(t));
                t = m;
            }
            if (!single_column_mm(t) || t->src[1] != mul || idx(t) < i) return 0;
            members.push_back(idx(t)); elided.push_back(idx(t));
            mm[s] = t;
        }
        // ---- shapes
        const int64_t D = kc->ne[0], n_kv = kc->ne[1], H_kv = kc->ne[2], H = ggml_nelements(q) / D;
        if (q->ne[0] != D || ggml_nelements(q) != mm[0]->ne[0] || mm[1]->ne[0] != D * H_kv || mm[2]->ne[0] != D * H_kv || H % H_kv != 0) return 0;
        // the attention phase gives one query head to one CTA and walks all n_kv positions: beyond ~1k positions the
        // split-KV kernels (b200_flash_attn_ext: vector / tensor-core decode) are faster even with the chain cut in two
        static const int64_t attn_max_kv = [] { const char * e = getenv("GGML_B200_CHAIN_ATTN_MAX_KV"); return e ? (int64_t) atoll(e) : (int64_t) 1024; }();
        if (n_kv > attn_max_kv) return 0;
        if (kc->ne[3] != 1 || vc->ne[0] != D || vc->ne[1] != n_kv || vc->ne[2] != H_kv || vc->ne[3] != 1) return 0;
        if (kc->data != k_leaf->data || vc->data != v_leaf->data || kc->nb[1] != k_leaf->nb[1] || vc->nb[1] != v_leaf->nb[1]) return 0;
        if (k_leaf->ne[0] != D * H_kv || v_leaf->ne[0] != D * H_kv) return 0;
        // the attention phase writes cache row `slot` at leaf->data + slot * leaf->nb[1]: the SET_ROWS it replaces must address
        // the cache the same way (no offset or merged view in between)
        if (sr_k->data != k_leaf->data || sr_k->nb[1] != k_leaf->nb[1] || sr_k->ne[0] != D * H_kv) return 0;
        if (sr_v->data != v_leaf->data || sr_v->nb[1] != v_leaf->nb[1] || sr_v->ne[0] != D * H_kv) return 0;
        for (const ggml_tensor * sr : {sr_k, sr_v}) {
            const ggml_tensor * ids = sr->src[1];
            if (ids->type != GGML_TYPE_I64 || ggml_nelements(ids) != 1 || !is_f32_vec(sr->src[0], D * H_kv) || sr->type != kc->type) return 0;
        }
        if (rq->src[2] || rk->src[2] || rq->src[1] != rk->src[1] || rq->src[1]->type != GGML_TYPE_I32 || ggml_nelements(rq->src[1]) != 1) return 0;
        if (memcmp(rq->op_params, rk->op_params, sizeof(rq->op_params)) != 0) return 0;
        if (rq->ne[0] != D || rk->ne[0] != D || !ggml_is_contiguous(rq) || !ggml_is_contiguous(rk)) return 0;
        const ggml_tensor * mask = fa->src[3];
        if (mask && (mask->type != GGML_TYPE_F16 || mask->ne[0] < n_kv)) return 0;
        if (!is_f32_vec(fa, D * H) && !(fa->ne[0] == D && ggml_nelements(fa) == D * H && ggml_is_contiguous(fa))) return 0;

        // ---- every other node inside [i, f] must be independent of what this block elides (e.g. the mask cast of layer 0)
        std::vector<int> foreign;
        {
            std::vector<char> in_block((size_t) (f - i + 1), 0);
            for (int m : members) in_block[(size_t) (m - i)] = 1;
            for (int j = i; j <= f; ++j) {
                if (in_block[(size_t) (j - i)]) continue;
                const ggml_tensor * n = g->nodes[j];
                for (int s = 0; s < GGML_MAX_SRC; ++s) {
                    const int k = n->src[s] ? idx(n->src[s]) : -1;
                    if (k >= i && k <= f && in_block[(size_t) (k - i)]) return 0;
                }
                if (ggml_is_empty(n) || (n->flags & GGML_TENSOR_FLAG_COMPUTE) == 0 || is_view_op(n->op) || n->op == GGML_OP_NONE) continue;   // nothing to run
                foreign.push_back(j);
            }
            // and the projections must not depend on a foreign node of the block
            for (int s = 0; s < 3; ++s) if (bias[s] && idx(bias[s]) >= i) return 0;
        }
        if (!uses_confined(members, elided)) return 0;

        // ---- phases
        plan.scratch_floats = std::max<int64_t>(plan.scratch_floats, D * (H + 2 * H_kv));
        const int64_t off[3] = {0, D * H, D * (H + H_kv)};
        pq.gemv.n_seg = 3; pq.gemv.glu_op = B200_GLU_NONE;
        for (int s = 0; s < 3; ++s) {
            b200_gemv_seg & sg = pq.gemv.seg[s];
            sg.w = b200_desc(mm[s]->src[0]);
            sg.bias = bias[s] ? keep(bias[s]) : nullptr;
            sg.dst = vec_desc((void *) (intptr_t) (off[s] * 4), mm[s]->ne[0]);       // scratch-relative; rebased at build time
        }
        if (!b200_gemv_fused_supported_rel(pq.gemv)) return 0;
        b200_phase pa{}; pa.kind = B200_PHASE_ATTN;
        b200_attn_phase & a = pa.attn;
        a.q = vec_desc((void *) (intptr_t) (off[0] * 4), D * H);
        a.k_new = vec_desc((void *) (intptr_t) (off[1] * 4), D * H_kv);
        a.v_new = vec_desc((void *) (intptr_t) (off[2] * 4), D * H_kv);
        a.k_cache = b200_desc(kc); a.v_cache = b200_desc(vc);
        a.pos = keep(rq->src[1]); a.slot = keep(sr_k->src[1]); a.slot_v = keep(sr_v->src[1]);
        a.mask = mask ? keep(mask) : nullptr;
        const int32_t * p = rq->op_params;
        a.rope.n_dims = p[1]; a.rope.mode = p[2]; a.rope.n_ctx_orig = p[4];
        memcpy(&a.rope.freq_base, p + 5, 4); memcpy(&a.rope.freq_scale, p + 6, 4); memcpy(&a.rope.ext_factor, p + 7, 4);
        memcpy(&a.rope.attn_factor, p + 8, 4); memcpy(&a.rope.beta_fast, p + 9, 4); memcpy(&a.rope.beta_slow, p + 10, 4);
        a.scale = scale;
        a.dst = vec_desc(fa->data, D * H);
        if (sinks) { if (sinks->ne[0] != H || ggml_nelements(sinks) != H) return 0; a.sinks = keep(sinks); }
        if ((a.rope.mode != 0 && a.rope.mode != 2) || a.rope.n_dims > D ||
            (D != 64 && D != 128) || kc->type != vc->type || (kc->type != GGML_TYPE_F16 && kc->type != GGML_TYPE_Q8_0)) return 0;

        // foreign nodes that only read graph inputs (the f32 -> f16 mask cast of the first layer) run BEFORE the projection
        // phase when what they write cannot disturb it, so that they do not split the chain
        std::vector<int> before, after;
        for (int j : foreign) {
            const ggml_tensor * n = g->nodes[j];
            bool early = after.empty() && !overlap(n->data, ggml_nbytes(n), pq.gemv.x.data, (size_t) pq.gemv.x.ne[0] * 4);
            for (int s = 0; s < GGML_MAX_SRC && early; ++s) early = !n->src[s] || n->src[s] == n || idx(n->src[s]) < i;      // (ggml_cast lists its own result as src[1])
            (early ? before : after).push_back(j);
        }
        b200_plan_item it{}; it.kind = B200_ITEM_PHASE; it.phase = pq; it.scratch_rel = true; it.first_node = i;
        for (int j : before) { b200_plan_item pn{}; pn.kind = B200_ITEM_NODE; pn.first_node = j; pn.n_nodes = 1; pn.hoisted = true; plan.items.push_back(pn); }
        plan.items.push_back(it);
        for (int j : after) { b200_plan_item pn{}; pn.kind = B200_ITEM_NODE; pn.first_node = j; pn.n_nodes = 1; pn.hoisted = true; plan.items.push_back(pn); }
        it.phase = pa; it.first_node = f;
        plan.items.push_back(it);
        return f - i + 1;
    }

    // ---------------------------------------------------------------- pattern G1: norm + matmul group
    int match_norm_group(int i) {
        b200_phase ph{}; ph.kind = B200_PHASE_GEMV;
        b200_gemv_fused_args & a = ph.gemv;
        // an OUTPUT-flagged norm (llama's result_norm) is still fused into the matmul phase; the norm itself is then
        // materialised by a plain kernel AFTER the phase (it is computed in place over x, which the phase reads)
        const ggml_tensor * mul = match_norm(i, a, true);
        if (!mul) return 0;
        const bool defer_norm = is_output(mul);
        std::vector<int> members = {i, i + 1}, elided = {i, i + 1};
        if (defer_norm) elided.clear();
        int j = i + 2;
        // gate / up / GLU
        if (j + 2 < g->n_nodes && single_column_mm(g->nodes[j]) && single_column_mm(g->nodes[j + 1]) && g->nodes[j + 2]->op == GGML_OP_GLU) {
            const ggml_tensor * m0 = g->nodes[j]; const ggml_tensor * m1 = g->nodes[j + 1]; const ggml_tensor * glu = g->nodes[j + 2];
            if (m0->src[1] == mul && m1->src[1] == mul && glu->src[0] && glu->src[1] && glu->op_params[1] == 0 &&
                ((glu->src[0] == m0 && glu->src[1] == m1) || (glu->src[0] == m1 && glu->src[1] == m0)) &&
                m0->ne[0] == m1->ne[0] && is_f32_vec(glu, m0->ne[0]) && m0->src[0]->type == m1->src[0]->type) {
                members.insert(members.end(), {j, j + 1, j + 2}); elided.insert(elided.end(), {j, j + 1});
                if (!uses_confined(members, elided)) return 0;
                const ggml_tensor * gate = glu->src[0]; const ggml_tensor * up = glu->src[1];
                a.n_seg = 2; a.glu_op = (int) ggml_get_glu_op(glu);
                memcpy(&a.glu_alpha, (const float *) glu->op_params + 2, 4); memcpy(&a.glu_limit, (const float *) glu->op_params + 3, 4);
                a.seg[0].w = b200_desc(gate->src[0]); a.seg[0].dst = b200_desc(glu);
                a.seg[1].w = b200_desc(up->src[0]);   a.seg[1].dst = b200_desc(glu);
                if (overlap(a.x.data, (size_t) a.x.ne[0] * 4, glu->data, ggml_nbytes(glu))) return 0;
                if (!b200_gemv_fused_supported(&a)) return 0;
                b200_plan_item it{}; it.kind = B200_ITEM_PHASE; it.phase = ph; it.first_node = i;
                plan.items.push_back(it);
                return 5;
            }
        }
        // n plain projections sharing the normalised activation
        const int n_mm = ggml_node_get_use_count(g, i + 1);
        if (n_mm < 1 || n_mm > B200_GEMV_MAX_SEG) return 0;
        a.n_seg = 0; a.glu_op = B200_GLU_NONE;
        int last = i + 1;
        for (; j < g->n_nodes && a.n_seg < n_mm; ++j) {
            const ggml_tensor * n = g->nodes[j];
            if (std::find(members.begin(), members.end(), j) != members.end()) continue;      // a bias ADD already taken
            if (!single_column_mm(n) || n->src[1] != mul) break;
            const ggml_tensor * bias = nullptr;
            const ggml_tensor * out = mm_with_bias(n, bias, members, elided);
            b200_gemv_seg & sg = a.seg[a.n_seg++];
            sg.w = b200_desc(n->src[0]); sg.bias = bias ? keep(bias) : nullptr; sg.dst = b200_desc(out);
            if (overlap(a.x.data, (size_t) a.x.ne[0] * 4, out->data, ggml_nbytes(out))) return 0;
            last = std::max(last, idx(out));
        }
        if (a.n_seg != n_mm) return 0;
        // the consumed nodes must be exactly i .. last (no foreign node in between)
        if ((int) members.size() != last - i + 1) return 0;
        if (!uses_confined(members, elided)) return 0;
        if (!b200_gemv_fused_supported(&a)) return 0;
        if (defer_norm) {
            // nothing the phase writes may alias the deferred norm's output, and nobody but the phase may read it before
            for (int s2 = 0; s2 < a.n_seg; ++s2)
                if (overlap(a.seg[s2].dst.data, (size_t) a.seg[s2].dst.ne[0] * 4, mul->data, ggml_nbytes(mul)) ||
                    overlap(a.seg[s2].dst.data, (size_t) a.seg[s2].dst.ne[0] * 4, g->nodes[i]->data, ggml_nbytes(g->nodes[i]))) return 0;
            if (ggml_node_get_use_count(g, i + 1) != n_mm) return 0;
        }
        b200_plan_item it{}; it.kind = B200_ITEM_PHASE; it.phase = ph; it.first_node = i;
        plan.items.push_back(it);
        if (defer_norm) { b200_plan_item pn{}; pn.kind = B200_ITEM_NODE; pn.first_node = i; pn.n_nodes = 2; pn.hoisted = true; plan.items.push_back(pn); }
        return last - i + 1;
    }

    // ---------------------------------------------------------------- pattern G2: matmul (+ add)
    int match_mm(int i) {
        const ggml_tensor * mm = g->nodes[i];
        if (!single_column_mm(mm)) return 0;
        b200_phase ph{}; ph.kind = B200_PHASE_GEMV;
        b200_gemv_fused_args & a = ph.gemv;
        a.x = read_desc(mm->src[1]); a.x.ne[1] = a.x.ne[2] = a.x.ne[3] = 1;
        a.n_seg = 1; a.glu_op = B200_GLU_NONE;
        const int64_t M = mm->ne[0];
        const ggml_tensor * out = mm; const ggml_tensor * addend = nullptr;
        int consumed = 1;
        std::vector<int> members = {i}, elided;
        if (ggml_node_get_use_count(g, i) == 1 && !is_output(mm)) {
            int j = i + 1;
            const ggml_tensor * cur = mm;
            std::vector<int> m2 = members, e2 = {i};
            // [GET_ROWS(cur)] [GET_ROWS(other)] ADD   (the GET_ROWS pair only on the last layer of a 1-token batch)
            if (j < g->n_nodes && is_identity_get_rows(g->nodes[j]) && g->nodes[j]->src[0] == cur && ggml_node_get_use_count(g, j) == 1) {
                cur = g->nodes[j]; m2.push_back(j); e2.push_back(j); ++j;
            }
            const ggml_tensor * other_copy = nullptr;
            if (j + 1 < g->n_nodes && is_identity_get_rows(g->nodes[j]) && g->nodes[j]->src[0] != cur && ggml_node_get_use_count(g, j) == 1 &&
                g->nodes[j + 1]->op == GGML_OP_ADD && (g->nodes[j + 1]->src[0] == g->nodes[j] || g->nodes[j + 1]->src[1] == g->nodes[j])) {
                other_copy = g->nodes[j]; m2.push_back(j); e2.push_back(j); ++j;
            }
            if (j < g->n_nodes && g->nodes[j]->op == GGML_OP_ADD) {
                const ggml_tensor * add = g->nodes[j];
                const ggml_tensor * other = add->src[0] == cur ? add->src[1] : add->src[1] == cur ? add->src[0] : nullptr;
                if (other && (!other_copy || other == other_copy) && is_f32_vec(add, M)) {
                    const ggml_tensor * src = other_copy ? other_copy->src[0] : other;
                    if (is_f32_vec(src, M)) {
                        m2.push_back(j);
                        if (uses_confined(m2, e2)) { out = add; addend = src; consumed = j - i + 1; members = m2; elided = e2; }
                    }
                }
            }
        }
        a.seg[0].w = b200_desc(mm->src[0]);
        const size_t xb = (size_t) a.x.ne[0] * 4, ob = (size_t) M * 4;
        // the residual may alias the output element for element (in-place ADD); any other overlap of what the phase
        // reads with what it writes is a hazard (ggml-alloc reuses dead buffers): fall back to the bare matmul
        void * out_data = out->data;
        if (addend && (overlap(a.x.data, xb, out->data, ob) || (addend->data != out->data && overlap(addend->data, ob, out->data, ob)))) {
            if (allow_redirect) { out_data = valloc((int64_t) ob); redirect[out] = out_data; }     // readers are checked at the end of the plan
            else { out = mm; addend = nullptr; consumed = 1; out_data = out->data; }
        }
        a.seg[0].dst = vec_desc(out_data, M);
        if (addend) { plan.pool.push_back(read_desc(addend)); a.seg[0].bias = &plan.pool.back(); }
        if (overlap(a.x.data, xb, out_data, ob)) return 0;
        if (!b200_gemv_fused_supported(&a)) return 0;
        b200_plan_item it{}; it.kind = B200_ITEM_PHASE; it.phase = ph; it.first_node = i;
        plan.items.push_back(it);
        return consumed;
    }
};

}  // namespace

// GEMV phase whose dst pointers are scratch-relative: validate with a stand-in base
bool b200_gemv_fused_supported_rel(const b200_gemv_fused_args & a) {
    b200_gemv_fused_args t = a;
    static float dummy[4];
    for (int s = 0; s < t.n_seg; ++s) t.seg[s].dst.data = (void *) (((uintptr_t) dummy + 15) & ~(uintptr_t) 15);
    return b200_gemv_fused_supported(&t) != 0;
}

static bool plan_build_once(ggml_cgraph * g, b200_plan & plan, bool allow_redirect) {
    plan.items.clear(); plan.pool.clear(); plan.scratch_floats = 0; plan.redirect_bytes = 0;
    matcher m(g, plan);
    m.allow_redirect = allow_redirect;
    for (int i = 0; i < g->n_nodes;) {
        ggml_tensor * n = g->nodes[i];
        if (ggml_is_empty(n) || (n->flags & GGML_TENSOR_FLAG_COMPUTE) == 0) { ++i; continue; }
        int c = 0;
        if (n->op == GGML_OP_RMS_NORM) { c = m.match_attention(i); if (!c) c = m.match_norm_group(i); }
        else if (n->op == GGML_OP_MUL_MAT) c = m.match_mm(i);
        if (c > 0) { i += c; continue; }
        b200_plan_item it{}; it.kind = B200_ITEM_NODE; it.first_node = i; it.n_nodes = 0;      // n_nodes filled by the executor (run_fused)
        plan.items.push_back(it);
        ++i;
    }
    const bool ok = m.redirects_ok();
    if (getenv("B200_PLAN_DEBUG")) {
        for (auto & kv : m.redirect) {
            auto u = m.redirect_uses.find(kv.first);
            fprintf(stderr, "plan: redirected %s uses %d phase-reads %d output %d\n", kv.first->name, m.uses(kv.first), u == m.redirect_uses.end() ? 0 : u->second, (int) m.is_output(kv.first));
        }
    }
    return ok;
}

void b200_plan_build(ggml_cgraph * g, b200_plan & plan) {
    if (!plan_build_once(g, plan, true)) plan_build_once(g, plan, false);      // a redirected tensor has a reader outside the chains
}
