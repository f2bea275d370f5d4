// gemv_stream.cu — K4s: streaming fused GEMV for decode (C-ABI b200_gemv_fused).
//
// Replaces, in ONE launch, the reference's per-matmul sequence
//   rms_norm (norm.cu:476) -> mul -> quantize_q8_1 (quantize.cu:4-48) -> mul_mat_vec_q (mmvq.cu:140-356) [-> add | glu]
// for up to 4 weight matrices that share one activation column.
//
// B200 design (see DESIGN.md "K4s"):
//   * persistent CTAs, one per SM; the rows of all segments are split across CTAs by BYTES, so every SM
//     streams the same amount of HBM;
//   * a producer warp moves weight rows HBM -> shared memory with 1-D TMA bulk copies
//     (cp.async.bulk ... mbarrier::complete_tx) into a ring of 3 x ~56 KB stages (measured: HBM efficiency of the
//     bulk copies needs >= ~32 KB per copy); alignment of the odd block formats (18/34/17/210-byte blocks)
//     is paid on shared memory reads, never on HBM transactions;
//   * 16 consumer warps do dp4a dot products out of shared memory against the q8_1 activation
//     (quantised once per CTA in the prologue; hoisted into registers per segment);
//   * a finalizer warp reduces the per-lane partial sums and applies the epilogue (bias/residual, GLU)
//     so consumers never wait on a shuffle chain or a CTA-wide barrier in the main loop;
//   * programmatic dependent launch: the producer starts prefetching weights before the previous kernel
//     in the stream has finished; only the activation read waits (griddepcontrol.wait);
//   * DECODE CHAIN (b200_chain_*): the same kernel runs a LIST of phases (every matmul of every layer plus an
//     attention phase with RoPE and the KV-cache write fused in) in ONE cooperative launch; phases are separated
//     by in-kernel grid barriers that only the consumers wait on — the producer keeps streaming the next phase's
//     weights into the ring, so HBM never drains at a matmul boundary.
#include "gemv_stream.cuh"

namespace b200 {

// the instantiations live in gemv_stream_i_*.cu
extern template int gs_launch<12, 12, 0>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<12, 12, 1>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<12, 12, 2>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<14, 14, 0>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<14, 14, 1>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<14, 14, 2>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<2, 2, 0>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<2, 2, 1>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<2, 2, 2>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<8, 8, 0>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<8, 8, 1>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<8, 8, 2>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<39, 39, 0>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<39, 39, 1>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<39, 39, 2>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<12, 14, 0>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<12, 14, 1>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<12, 14, 2>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<2, 14, 0>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<2, 14, 1>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<2, 14, 2>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<2, 8, 0>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<2, 8, 1>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<2, 8, 2>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<8, 39, 0>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<8, 39, 1>(const gs_params &, size_t, int, cudaStream_t);
extern template int gs_launch<8, 39, 2>(const gs_params &, size_t, int, cudaStream_t);

// ---------------------------------------------------------------- host side
static bool gs_type_ok(int t) { return t == B200_Q4_0 || t == B200_Q8_0 || t == B200_MXFP4 || t == B200_Q4_K || t == B200_Q6_K; }
static int  gs_wpi(int t) { return (t == B200_Q4_K || t == B200_Q6_K) ? 64 : 32; }

// validate one GEMV phase and fill everything that does not depend on the ring configuration
static bool gs_build_phase(const b200_gemv_fused_args * a, gs_phase & ph, int64_t & need, int & act_qs, int types[2]) {
    if (!a || a->n_seg < 1 || a->n_seg > B200_GEMV_MAX_SEG) return false;
    const b200_tensor & x = a->x;
    const int64_t K = x.ne[0];
    if (x.type != B200_F32 || x.nb[0] != 4 || x.ne[1] * x.ne[2] * x.ne[3] != 1 || K < 32 || K % 32 != 0 || K > 65536) return false;
    if (((uintptr_t) x.data) % 16 != 0) return false;
    if (a->norm_w && (a->norm_w->type != B200_F32 || a->norm_w->nb[0] != 4 || a->norm_w->ne[0] != K ||
                      nelements(*a->norm_w) != K || ((uintptr_t) a->norm_w->data) % 16 != 0)) return false;
    const bool glu = a->glu_op != B200_GLU_NONE;
    if (glu && (a->n_seg != 2 || a->glu_op < B200_GLU_REGLU || a->glu_op > B200_GLU_GEGLU_QUICK)) return false;
    for (int s = 0; s < a->n_seg; ++s) {
        const b200_gemv_seg & g = a->seg[s];
        const b200_tensor & w = g.w;
        if (!gs_type_ok(w.type) || w.ne[0] != K || w.ne[1] < 1 || w.ne[2] != 1 || w.ne[3] != 1) return false;
        if (K % b200_blck_size(w.type) != 0) return false;
        if (w.nb[0] != (int64_t) b200_type_size(w.type) || w.nb[1] != (int64_t) b200_row_size(w.type, K)) return false;
        if (((uintptr_t) w.data) % 16 != 0 || w.ne[1] > (1 << 30)) return false;
        if (!(glu && s == 0)) {
            const b200_tensor & d = g.dst;
            if (d.type != B200_F32 || d.nb[0] != 4 || d.ne[0] != w.ne[1] || nelements(d) != w.ne[1]) return false;
        }
        if (g.bias && (g.bias->type != B200_F32 || g.bias->nb[0] != 4 || g.bias->ne[0] != w.ne[1] || nelements(*g.bias) != w.ne[1])) return false;
        if (types[0] < 0 || types[0] == w.type) types[0] = w.type;
        else if (types[1] < 0 || types[1] == w.type) types[1] = w.type;
        else return false;
        need = std::max<int64_t>(need, w.nb[1] * (glu ? 2 : 1) + 64);
    }
    if (glu) {
        const b200_tensor & g = a->seg[0].w; const b200_tensor & u = a->seg[1].w;
        if (g.type != u.type || g.ne[1] != u.ne[1] || a->seg[0].bias) return false;
    }
    act_qs = std::max(act_qs, (int) ((K + (K / 128 + 1) * 16 + 15) & ~15));

    memset(&ph, 0, sizeof(ph));
    ph.kind = 0;
    ph.K = (int32_t) K; ph.x = (const float *) x.data; ph.norm_w = a->norm_w ? (const float *) a->norm_w->data : nullptr; ph.eps = a->eps;
    ph.glu_op = a->glu_op; ph.glu_alpha = a->glu_alpha; ph.glu_limit = a->glu_limit;
    ph.xq_out = (blk_q8_1 *) a->xq_out;
    ph.n_ws = glu ? 1 : a->n_seg;
    for (int s = 0; s < ph.n_ws; ++s) {
        gs_seg & g = ph.seg[s];
        const b200_gemv_seg & src = glu ? a->seg[1] : a->seg[s];
        g.nsub = glu ? 2 : 1;
        g.w[0] = (const uint8_t *) (glu ? a->seg[0].w.data : src.w.data);
        g.w[1] = glu ? (const uint8_t *) a->seg[1].w.data : nullptr;
        g.bias = src.bias ? (const float *) src.bias->data : nullptr;
        g.dst = (float *) src.dst.data;
        g.M = (int32_t) src.w.ne[1]; g.rb = (int32_t) src.w.nb[1]; g.type = src.w.type;
        g.ipr = (int32_t) (K / gs_wpi(g.type)); g.spr = (g.ipr + 31) / 32;
        // the fewest warps per row whose lanes can hold the activations of their K slice in registers: every extra
        // K split costs a cross-warp reduction per pass
        const int max_spw = gs_wpi(g.type) == 64 ? 2 : 4;
        int KS = 1; while (KS < GS_NW && (g.spr + KS - 1) / KS > max_spw) KS *= 2;
        const int force = (a->flags >> 16) & 0xF;                       // tuning: log2(KS) + 1
        if (force) KS = std::max(KS, std::min(1 << (force - 1), GS_NW));
        int spw = (g.spr + KS - 1) / KS;
        if (spw == 3) spw = 4;
        if (spw > max_spw) return false;                                // row too long for the register-resident activation
        g.KS = KS; g.spw = spw;
    }
    return true;
}

// rows per chunk once the stage payload is known
static bool gs_finish_phase(gs_phase & ph, int stage_payload) {
    int64_t prefix = 0;
    for (int s = 0; s < ph.n_ws; ++s) {
        gs_seg & g = ph.seg[s];
        int R = (stage_payload - 64) / (g.rb * g.nsub);
        if (R > 64) R = 64;
        if (R < 1) return false;
        const int opp = (g.KS == 1 ? 4 : GS_RB) / g.nsub;  // output rows per warp pass (gemv_stream.cuh, consume_segment)
        if (R > opp) R -= R % opp;                         // whole warp passes per chunk
        g.R = R;
        g.byte_prefix = prefix;
        prefix += (int64_t) g.rb * g.nsub * g.M;
    }
    ph.total_bytes = prefix;
    return true;
}

struct gs_config { int n_stages, stage_payload, act_qs, attn_floats; size_t smem; };

static bool gs_make_config(int flags, int64_t need, int act_qs, int attn_floats, gs_config & cfg) {
    cfg.n_stages = ((flags >> 4) & 0xF) ? std::min((flags >> 4) & 0xF, GS_MAX_STAGES) : GS_STAGES;
    const int64_t want = ((flags >> 8) & 0xFF) ? (int64_t) ((flags >> 8) & 0xFF) * 1024 : GS_STAGE_BYTES;
    const int64_t fixed = act_qs + act_qs / 4 + 64 + (int64_t) attn_floats * 4 + 64;
    int64_t payload = std::max<int64_t>(want, need);
    auto total_for = [&](int64_t pl) { return (int64_t) cfg.n_stages * (((pl + 127) & ~127) + GS_STAGE_SLACK) + fixed; };
    if (total_for(payload) > GS_SMEM_MAX) {            // shrink the stages (never below one row) to fit
        const int64_t fit = ((GS_SMEM_MAX - fixed) / cfg.n_stages - GS_STAGE_SLACK) & ~127;
        if (fit < need) return false;
        payload = fit;
    }
    cfg.stage_payload = (int) ((payload + 127) & ~127);
    cfg.act_qs = act_qs; cfg.attn_floats = attn_floats;
    cfg.smem = (size_t) total_for(cfg.stage_payload);
    return true;
}


template <int MODE>
static int gs_launch_types(const gs_params & p, size_t smem, int grid, cudaStream_t st, int t0, int t1) {
    if (t1 < 0) t1 = t0;
    if (t0 == t1) {
        switch (t0) {
            case B200_Q4_K:  return gs_launch<B200_Q4_K, B200_Q4_K, MODE>(p, smem, grid, st);
            case B200_Q6_K:  return gs_launch<B200_Q6_K, B200_Q6_K, MODE>(p, smem, grid, st);
            case B200_Q4_0:  return gs_launch<B200_Q4_0, B200_Q4_0, MODE>(p, smem, grid, st);
            case B200_Q8_0:  return gs_launch<B200_Q8_0, B200_Q8_0, MODE>(p, smem, grid, st);
            case B200_MXFP4: return gs_launch<B200_MXFP4, B200_MXFP4, MODE>(p, smem, grid, st);
        }
    }
    const int lo = t0 < t1 ? t0 : t1, hi = t0 < t1 ? t1 : t0;
    if (lo == B200_Q4_K && hi == B200_Q6_K)  return gs_launch<B200_Q4_K, B200_Q6_K, MODE>(p, smem, grid, st);
    if (lo == B200_Q4_0 && hi == B200_Q6_K)  return gs_launch<B200_Q4_0, B200_Q6_K, MODE>(p, smem, grid, st);
    if (lo == B200_Q4_0 && hi == B200_Q8_0)  return gs_launch<B200_Q4_0, B200_Q8_0, MODE>(p, smem, grid, st);
    if (lo == B200_Q8_0 && hi == B200_MXFP4) return gs_launch<B200_Q8_0, B200_MXFP4, MODE>(p, smem, grid, st);
    set_error("b200 streaming GEMV: unsupported type pair (%d, %d)", t0, t1);
    return B200_ERR_UNSUPPORTED;
}

static int gs_sm_count() {
    static int n[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (!n[dev]) { if (cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n[dev] <= 0) n[dev] = 148; }
    return n[dev];
}

static bool gs_build_single(const b200_gemv_fused_args * a, gs_params & p, size_t & smem, int types[2]) {
    int64_t need = 0; int act_qs = 0;
    types[0] = types[1] = -1;
    memset(&p, 0, sizeof(p));
    if (!gs_build_phase(a, p.inl, need, act_qs, types)) return false;
    gs_config cfg;
    if (!gs_make_config(a->flags, need, act_qs, 0, cfg)) return false;
    if (!gs_finish_phase(p.inl, cfg.stage_payload)) return false;
    p.n_phases = 1; p.act_qs_bytes = cfg.act_qs; p.stage_stride = cfg.stage_payload + GS_STAGE_SLACK;
    p.n_stages = cfg.n_stages; p.flags = a->flags; p.attn_floats = 0;
    smem = cfg.smem;
    return true;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_gemv_fused_supported(const b200_gemv_fused_args * a) {
    gs_params p; size_t smem; int types[2];
    return gs_build_single(a, p, smem, types) ? 1 : 0;
}

extern "C" int b200_gemv_fused(const b200_gemv_fused_args * a, void * stream) {
    gs_params p; size_t smem; int types[2];
    if (!gs_build_single(a, p, smem, types)) { set_error("b200_gemv_fused: unsupported arguments"); return B200_ERR_UNSUPPORTED; }
    if (stream == B200_STREAM_VALIDATE) return B200_OK;
    int64_t rows = 0;
    for (int s = 0; s < p.inl.n_ws; ++s) rows += p.inl.seg[s].M;
    int grid = a->n_sm > 0 ? a->n_sm : gs_sm_count();
    if (grid > rows) grid = (int) rows;                // never launch more CTAs than there are rows
    return gs_launch_types<0>(p, smem, grid, (cudaStream_t) stream, types[0], types[1]);
}

// ---------------------------------------------------------------- decode chain
struct b200_chain {
    gs_params p;
    size_t smem = 0;
    int grid = 0, t0 = -1, t1 = -1, device = 0;
    gs_phase * d_phases = nullptr;
    unsigned int * d_ctl = nullptr;    // [0] launches completed (the epoch of the tags), [1] CTAs that have left the current launch
    unsigned long long * d_shadow = nullptr;   // tagged words of every output a later phase reads (private to this chain)
    int2 * d_ranges = nullptr;         // rows of every (phase, CTA, segment)
    int n_gemv = 0, n_attn = 0;
    unsigned long long * d_dbg = nullptr;
    bool peer = false;                 // row split over peer memory
};

static bool gs_build_attn(const b200_attn_phase * a, gs_phase & ph, int & attn_floats) {
    const b200_tensor & q = a->q; const b200_tensor & k = a->k_cache; const b200_tensor & v = a->v_cache;
    const int64_t D = k.ne[0], n_kv = k.ne[1], H_kv = k.ne[2];
    if (D != 64 && D != 128) return false;
    if (k.type != v.type || (k.type != B200_F16 && k.type != B200_Q8_0) || k.ne[3] != 1 || v.ne[3] != 1) return false;
    if (k.nb[0] != (int64_t) b200_type_size(k.type) || v.nb[0] != k.nb[0]) return false;
    if (v.ne[0] != D || v.ne[1] != n_kv || v.ne[2] != H_kv || n_kv < 1 || H_kv < 1) return false;
    // f16 rows are read as 16-byte words; q8_0 rows (34-byte blocks) as aligned words around 2-byte aligned block starts
    const int64_t al = k.type == B200_F16 ? 16 : 2;
    if (((uintptr_t) k.data | (uintptr_t) v.data) % al || k.nb[1] % al || k.nb[2] % al || v.nb[1] % al || v.nb[2] % al) return false;
    if (q.type != B200_F32 || q.nb[0] != 4 || nelements(q) % D != 0) return false;
    const int64_t H = nelements(q) / D;
    if (H < 1 || H % H_kv != 0) return false;
    if (a->k_new.type != B200_F32 || a->v_new.type != B200_F32 || nelements(a->k_new) != D * H_kv || nelements(a->v_new) != D * H_kv) return false;
    if (a->dst.type != B200_F32 || nelements(a->dst) != D * H) return false;
    if (!a->pos || a->pos->type != B200_I32 || !a->slot || a->slot->type != B200_I64 || (a->slot_v && a->slot_v->type != B200_I64)) return false;
    if (a->mask && (a->mask->type != B200_F16 || a->mask->nb[0] != 2 || a->mask->ne[0] < n_kv)) return false;
    const b200_rope_params & r = a->rope;
    if ((r.mode != 0 && r.mode != 2) || r.n_dims <= 0 || r.n_dims % 2 || r.n_dims > D) return false;
    if (r.ext_factor != 0.0f && (r.freq_scale <= 0.0f || r.freq_base <= 1.0f || r.n_ctx_orig <= 0)) return false;
    if (a->sinks && (a->sinks->type != B200_F32 || a->sinks->nb[0] != 4 || a->sinks->ne[0] != H)) return false;
    memset(&ph, 0, sizeof(ph));
    ph.kind = 1;
    gs_attn & g = ph.attn;
    g.q = (const float *) q.data; g.k_new = (const float *) a->k_new.data; g.v_new = (const float *) a->v_new.data;
    g.k_cache = (char *) k.data; g.v_cache = (char *) v.data; g.mask = a->mask ? (const __half *) a->mask->data : nullptr;
    g.dst = (float *) a->dst.data; g.pos = (const int32_t *) a->pos->data; g.slot = (const int64_t *) a->slot->data;
    g.slot_v = a->slot_v ? (const int64_t *) a->slot_v->data : g.slot;
    g.k_nb1 = k.nb[1]; g.k_nb2 = k.nb[2]; g.v_nb1 = v.nb[1]; g.v_nb2 = v.nb[2];
    g.D = (int32_t) D; g.H = (int32_t) H; g.H_kv = (int32_t) H_kv; g.n_kv = (int32_t) n_kv;
    g.rope_mode = r.mode; g.n_dims = r.n_dims; g.scale = a->scale;
    g.theta_scale = powf(r.freq_base, -2.0f / (float) r.n_dims); g.freq_scale = r.freq_scale;
    g.ext_factor = r.ext_factor; g.attn_factor = r.attn_factor;
    {   // ggml_rope_yarn_corr_dims (reference ggml/src/ggml.c)
        const float PI = 3.14159265358979323846f;
        auto corr_dim = [&](float n_rot) { return r.n_dims * logf(r.n_ctx_orig / (n_rot * 2.0f * PI)) / (2.0f * logf(r.freq_base)); };
        g.corr0 = r.ext_factor != 0.0f ? fmaxf(0.0f, floorf(corr_dim(r.beta_fast))) : 0.0f;
        g.corr1 = r.ext_factor != 0.0f ? fminf((float) r.n_dims - 1.0f, ceilf(corr_dim(r.beta_slow))) : 1.0f;
    }
    g.sinks = a->sinks ? (const float *) a->sinks->data : nullptr;
    g.kv_type = (uint32_t) k.type;
    attn_floats = std::max(attn_floats, (int) (3 * D + GS_NW * (D + 2)));
    return true;
}

// ---- who produced what.  Every value a phase reads from an earlier phase of the chain travels as a tagged word in a shadow
// array (gemv_stream.cuh, gs_seg).  An OUTPUT of phase j is a run of f32: a GEMV segment's dst or an attention phase's dst.
struct gs_output { const char * p0; const char * p1; int phase; int seg; unsigned long long * ll; bool consumed; };

// the most recent output before phase i that overlaps the f32 range [ptr, ptr + n); null: plain memory (written before the launch)
static gs_output * gs_find_output(std::vector<gs_output> & outs, int i, const void * ptr, int64_t n) {
    const char * p0 = (const char *) ptr, * p1 = p0 + 4 * n;
    gs_output * best = nullptr;
    for (gs_output & o : outs) {
        // an EMPTY output (row split: a rank that owns no rows of a matrix still names where its slice would start, an address
        // inside the full dst) counts as produced: the other ranks' rows of that phase fill the range
        const bool hit = o.p0 == o.p1 ? (o.p0 >= p0 && o.p0 < p1) : (o.p0 < p1 && o.p1 > p0);
        if (o.phase < i && hit && (!best || o.phase > best->phase)) best = &o;
    }
    return best;
}

static b200_chain * gs_chain_create(const b200_phase * phases_in, int n_phases, int flags, const b200_peer_group * pg) {
    if (!phases_in || n_phases < 1 || n_phases > GS_MAX_PHASES) { set_error("b200_chain_create: bad phase count"); return nullptr; }
    std::vector<b200_phase> phv(phases_in, phases_in + n_phases);        // private copy: zero-row segments are removed below
    uintptr_t a0 = 0, a1 = 0, act1 = 0;
    if (pg) {
        if (pg->n_ranks < 1 || pg->n_ranks > B200_MAX_RANKS || pg->rank < 0 || pg->rank >= pg->n_ranks || pg->delta[pg->rank] != 0 ||
            pg->total_ctas < 1 || !pg->arena || !pg->sync) { set_error("b200_chain_create_peer: bad peer group"); return nullptr; }
        a0 = (uintptr_t) pg->arena; a1 = a0 + pg->arena_bytes; const uintptr_t s0 = (uintptr_t) pg->sync;
        if (pg->act_bytes == 0 || pg->act_bytes % 16 || pg->ll_offset % 16 || pg->ll_offset < pg->act_bytes || pg->ll_offset + 2 * pg->act_bytes > pg->arena_bytes) {
            set_error("b200_chain_create_peer: bad arena layout (act_bytes / ll_offset)"); return nullptr;
        }
        act1 = a0 + pg->act_bytes;
        if (s0 < a0 || s0 + sizeof(unsigned int) * (size_t) n_phases > a1 || s0 % 4 ||
            (s0 < a0 + pg->ll_offset + 2 * pg->act_bytes && s0 + sizeof(unsigned int) * (size_t) n_phases > a0 + pg->ll_offset)) {
            set_error("b200_chain_create_peer: sync counters outside the arena (or inside the tagged shadow)"); return nullptr;
        }
        for (int i = 0; i < n_phases; ++i) {
            if (phv[i].kind != B200_PHASE_GEMV) continue;
            const b200_gemv_fused_args & g = phv[i].gemv;
            if (g.n_seg < 1 || g.n_seg > B200_GEMV_MAX_SEG) { set_error("b200_chain_create_peer: phase %d: bad segment count", i); return nullptr; }
            for (int k = 0; k < g.n_seg; ++k) {
                if (g.glu_op != B200_GLU_NONE && k == 0) continue;
                const uintptr_t d0 = (uintptr_t) g.seg[k].dst.data, d1 = d0 + 4 * (size_t) g.seg[k].w.ne[1];
                if (d0 < a0 || d1 > act1) { set_error("b200_chain_create_peer: phase %d segment %d: dst outside the symmetric arena", i, k); return nullptr; }
            }
        }
    }
    // ---- outputs (before empty slices are dropped: they still say which phase fills a range)
    std::vector<gs_output> outs;
    for (int i = 0; i < n_phases; ++i) {
        if (phv[i].kind == B200_PHASE_GEMV) {
            const b200_gemv_fused_args & g = phv[i].gemv;
            for (int k = 0; k < g.n_seg && k < B200_GEMV_MAX_SEG; ++k) {
                if (g.glu_op != B200_GLU_NONE && k == 0) continue;                 // the gate segment has no dst of its own
                const char * d0 = (const char *) g.seg[k].dst.data;
                if (d0) outs.push_back({d0, d0 + 4 * (size_t) std::max<int64_t>(g.seg[k].w.ne[1], 0), i, 0, nullptr, false});
            }
        } else if (phv[i].kind == B200_PHASE_ATTN) {
            const b200_tensor & d = phv[i].attn.dst;
            outs.push_back({(const char *) d.data, (const char *) d.data + 4 * (size_t) nelements(d), i, -1, nullptr, false});
        }
    }
    if (pg) {
        // a rank may own no rows of a small matrix (split boundaries round down): the segment only described the slice
        for (int i = 0; i < n_phases; ++i) {
            if (phv[i].kind != B200_PHASE_GEMV) continue;
            b200_gemv_fused_args & g = phv[i].gemv;
            if (g.glu_op != B200_GLU_NONE) { if (g.seg[1].w.ne[1] == 0) phv[i].kind = B200_PHASE_NOP; continue; }
            int n = 0;
            for (int k = 0; k < g.n_seg; ++k) if (g.seg[k].w.ne[1] > 0) g.seg[n++] = g.seg[k];
            g.n_seg = n;
            if (n == 0) phv[i].kind = B200_PHASE_NOP;
        }
    }
    const b200_phase * phases = phv.data();
    std::vector<gs_phase> ph((size_t) n_phases);
    int64_t need = 0; int act_qs = 64, attn_floats = 0; int types[2] = {-1, -1};
    b200_chain * ch = new b200_chain;
    for (int i = 0; i < n_phases; ++i) {
        bool ok;
        if (phases[i].kind == B200_PHASE_GEMV) { ok = gs_build_phase(&phases[i].gemv, ph[i], need, act_qs, types); ch->n_gemv++; }
        else if (phases[i].kind == B200_PHASE_ATTN) { ok = gs_build_attn(&phases[i].attn, ph[i], attn_floats); ch->n_attn++; }
        else if (phases[i].kind == B200_PHASE_NOP)  { memset(&ph[i], 0, sizeof(gs_phase)); ph[i].kind = 2; ok = true; }
        else ok = false;
        if (!ok) { set_error("b200_chain_create: phase %d is not supported", i); delete ch; return nullptr; }
        ph[i].self_tag = (uint32_t) i;
    }
    // ---- outputs and their shadows.  Row split: the shadow of an arena address is arena-mapped (every rank's copy sits at the
    // same offset, so a peer's copy is one delta away) and a consumer may read rows other ranks produced in the same phase of
    // THEIR chain; everything else gets a private slot in this chain's own shadow pool.
    size_t pool_words = 0;
    auto in_arena = [&](const void * p, int64_t n) { return pg && (uintptr_t) p >= a0 && (uintptr_t) p + 4 * (size_t) n <= act1; };
    auto arena_ll = [&](const void * p) { return (unsigned long long *) ((char *) pg->arena + pg->ll_offset + 2 * ((uintptr_t) p - a0)); };
    for (gs_output & o : outs) {
        const int64_t n = (o.p1 - o.p0) / 4;
        if (in_arena(o.p0, n)) o.ll = arena_ll(o.p0);
        else { o.ll = (unsigned long long *) (uintptr_t) (8 * pool_words + 8); pool_words += (size_t) ((n + 1) & ~(int64_t) 1); }     // offset + 8 for now (0 = none); 16 B aligned slots
    }
    // ---- inputs: x, epilogue operands, q / k / v
    struct use { const void * ptr; int64_t n; const unsigned long long ** ll; uint32_t * tag; };
    bool bad = false;
    auto resolve = [&](int i, const void * ptr, int64_t n, const unsigned long long *& ll, uint32_t & tag, const char * what) {
        ll = nullptr; tag = 0;
        if (!ptr) return;
        gs_output * o = gs_find_output(outs, i, ptr, n);
        if (!o) return;                                            // plain memory written before the launch
        const char * p0 = (const char *) ptr;
        if (in_arena(ptr, n)) {
            // row split: [ptr, ptr + n) may be the concatenation of every rank's slice of phase o->phase; all of it carries that tag
            if (!in_arena(o->p0, (o->p1 - o->p0) / 4)) { set_error("b200_chain_create_peer: phase %d: %s mixes arena and private memory", i, what); bad = true; return; }
            ll = arena_ll(ptr); tag = (uint32_t) o->phase; o->consumed = true;
            // other outputs of the same phase that overlap the range are consumed too (q, k, v of one projection phase)
            for (gs_output & q : outs) if (q.phase == o->phase && q.p0 < p0 + 4 * n && q.p1 > p0) q.consumed = true;
            return;
        }
        if (p0 < o->p0 || p0 + 4 * n > o->p1) { set_error("b200_chain_create: phase %d: %s is only partly produced by phase %d", i, what, o->phase); bad = true; return; }
        ll = (const unsigned long long *) ((uintptr_t) o->ll + 2 * (size_t) (p0 - o->p0));      // still pool-relative (+8)
        tag = (uint32_t) o->phase; o->consumed = true;
    };
    for (int i = 0; i < n_phases && !bad; ++i) {
        if (ph[i].kind == 0) {
            resolve(i, ph[i].x, ph[i].K, ph[i].x_ll, ph[i].x_tag, "x");
            // the prologue reads the shadow of x two words (16 bytes) at a time
            const uintptr_t xa = (uintptr_t) ph[i].x_ll - (in_arena(ph[i].x, ph[i].K) ? 0 : 8);
            if (ph[i].x_ll && xa % 16) { set_error("b200_chain_create: phase %d: x is not 8-byte aligned inside its producer's dst", i); bad = true; }
            for (int s = 0; s < ph[i].n_ws && !bad; ++s) {
                gs_seg & g = ph[i].seg[s];
                resolve(i, g.bias, g.M, g.bias_ll, g.bias_tag, "bias");
            }
        } else if (ph[i].kind == 1) {
            gs_attn & g = ph[i].attn;
            resolve(i, g.q, (int64_t) g.D * g.H, g.q_ll, g.q_tag, "q");
            resolve(i, g.k_new, (int64_t) g.D * g.H_kv, g.k_ll, g.k_tag, "k");
            resolve(i, g.v_new, (int64_t) g.D * g.H_kv, g.v_ll, g.v_tag, "v");
        }
    }
    if (bad) { delete ch; return nullptr; }
    int n_final = 0, final_phase[4] = {0, 0, 0, 0};
    if (pg) {
        for (int i = 0; i < n_phases; ++i) {
            if (ph[i].kind != 0) continue;
            bool any = false;
            for (gs_output & o : outs) if (o.phase == i && o.consumed) any = true;
            ph[i].peer_mode = any ? 1 : 2;
            if (!any) {
                if (n_final == 4) { set_error("b200_chain_create_peer: more than 4 phases whose outputs leave the chain"); delete ch; return nullptr; }
                final_phase[n_final++] = i;
            }
        }
    }
    gs_config cfg;
    if (types[0] < 0 || !gs_make_config(flags, need, act_qs, attn_floats, cfg)) { set_error("b200_chain_create: no ring configuration fits"); delete ch; return nullptr; }
    for (int i = 0; i < n_phases; ++i)
        if (ph[i].kind == 0 && !gs_finish_phase(ph[i], cfg.stage_payload)) { set_error("b200_chain_create: phase %d: row does not fit a stage", i); delete ch; return nullptr; }
    cudaGetDevice(&ch->device);
    ch->peer = pg != nullptr;
    const size_t pool_bytes = 8 * pool_words + 64;
    if (cudaMalloc((void **) &ch->d_phases, sizeof(gs_phase) * (size_t) n_phases) != cudaSuccess ||
        cudaMalloc((void **) &ch->d_ctl, 2 * sizeof(unsigned int)) != cudaSuccess || cudaMemset(ch->d_ctl, 0, 2 * sizeof(unsigned int)) != cudaSuccess ||
        cudaMalloc((void **) &ch->d_shadow, pool_bytes) != cudaSuccess || cudaMemset(ch->d_shadow, 0, pool_bytes) != cudaSuccess) {
        set_error("b200_chain_create: device allocation failed"); cudaGetLastError();
        if (ch->d_phases) cudaFree(ch->d_phases); if (ch->d_ctl) cudaFree(ch->d_ctl); if (ch->d_shadow) cudaFree(ch->d_shadow);
        delete ch; return nullptr;
    }
    // pool-relative shadow addresses -> device addresses; outputs nobody reads get no shadow
    auto fix = [&](const unsigned long long *& ll, const void * ptr, int64_t n) {
        if (ll && !in_arena(ptr, n)) ll = (const unsigned long long *) ((char *) ch->d_shadow + ((uintptr_t) ll - 8));
    };
    for (gs_output & o : outs) {
        unsigned long long * ll = nullptr;
        if (o.consumed) ll = in_arena(o.p0, (o.p1 - o.p0) / 4) ? o.ll : (unsigned long long *) ((char *) ch->d_shadow + ((uintptr_t) o.ll - 8));
        if (o.seg < 0) { ph[o.phase].attn.dst_ll = ll; continue; }
        if (ph[o.phase].kind != 0) continue;                       // dropped: this rank owns no rows of it
        for (int s2 = 0; s2 < ph[o.phase].n_ws; ++s2)
            if ((const char *) ph[o.phase].seg[s2].dst == o.p0 && o.p1 > o.p0) ph[o.phase].seg[s2].dst_ll = ll;
    }
    for (int i = 0; i < n_phases; ++i) {
        if (ph[i].kind == 0) {
            fix(ph[i].x_ll, ph[i].x, ph[i].K);
            for (int s = 0; s < ph[i].n_ws; ++s) fix(ph[i].seg[s].bias_ll, ph[i].seg[s].bias, ph[i].seg[s].M);
        } else if (ph[i].kind == 1) {
            gs_attn & g = ph[i].attn;
            fix(g.q_ll, g.q, (int64_t) g.D * g.H); fix(g.k_ll, g.k_new, (int64_t) g.D * g.H_kv); fix(g.v_ll, g.v_new, (int64_t) g.D * g.H_kv);
        }
    }
    // rows of every segment for every CTA: bytes of a phase are split evenly, a CTA takes the rows that START in its share
    const int grid = gs_sm_count();
    std::vector<int2> ranges((size_t) n_phases * grid * B200_GEMV_MAX_SEG, make_int2(0, 0));
    // Tuning (flags bits 24-27: s1 in steps of 5 %, bits 28-30: s2 in steps of 1 %): the CTAs that ran an attention phase come
    // back to the GEMV code with a cold instruction cache and finish the next phases late (profiles/r02b_timeline_ctas_final.txt);
    // they get (1 - s1) of an even share in the phase right after attention and (1 - s2) in the one after that.
    // Default 20 % / 4 % (same-box A/B on the 8B chain: 567.1 -> 571.1 tok/s; 20-40 % / 4-7 % all within 0.1 %); bits 24-27 = 0xF: even split.
    const int f1 = (flags >> 24) & 0xF, f2 = (flags >> 28) & 0x7;
    const bool even = f1 == 0xF || (pg && !f1 && !f2);       // row split over peer memory: even unless asked for (not measured there)
    const double skew1 = even ? 0.0 : (f1 || f2) ? 0.05 * f1 : 0.20, skew2 = even ? 0.0 : (f1 || f2) ? 0.01 * f2 : 0.04;
    std::vector<int64_t> bound((size_t) grid + 1);
    for (int i = 0; i < n_phases; ++i) {
        if (ph[i].kind != 0) continue;
        int n_attn = 0; double sk = 0.0;
        if (i >= 1 && ph[i - 1].kind == 1)      { n_attn = std::min(ph[i - 1].attn.H, grid); sk = skew1; }
        else if (i >= 2 && ph[i - 2].kind == 1) { n_attn = std::min(ph[i - 2].attn.H, grid); sk = skew2; }
        {
            double tot_w = 0.0, acc_w = 0.0;
            for (int c = 0; c < grid; ++c) tot_w += c < n_attn ? 1.0 - sk : 1.0;
            bound[0] = 0;
            for (int c = 0; c < grid; ++c) { acc_w += c < n_attn ? 1.0 - sk : 1.0; bound[(size_t) c + 1] = (int64_t) ((double) ph[i].total_bytes * (acc_w / tot_w)); }
            bound[(size_t) grid] = ph[i].total_bytes;
            if (sk == 0.0) for (int c = 0; c <= grid; ++c) bound[(size_t) c] = ph[i].total_bytes * (int64_t) c / grid;      // exact integer split
        }
        for (int c = 0; c < grid; ++c) {
            const int64_t lo = bound[(size_t) c], hi = bound[(size_t) c + 1];
            for (int s2 = 0; s2 < ph[i].n_ws; ++s2) {
                const gs_seg & g = ph[i].seg[s2];
                const int64_t rbw = (int64_t) g.rb * g.nsub, tot = rbw * g.M;
                const int64_t a = std::min(std::max(lo - g.byte_prefix, (int64_t) 0), tot), b = std::min(std::max(hi - g.byte_prefix, (int64_t) 0), tot);
                ranges[((size_t) i * grid + c) * B200_GEMV_MAX_SEG + s2] = make_int2((int) ((a + rbw - 1) / rbw), (int) ((b + rbw - 1) / rbw));
            }
        }
    }
    if (cudaMalloc((void **) &ch->d_ranges, sizeof(int2) * ranges.size()) != cudaSuccess ||
        cudaMemcpy(ch->d_ranges, ranges.data(), sizeof(int2) * ranges.size(), cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMemcpy(ch->d_phases, ph.data(), sizeof(gs_phase) * (size_t) n_phases, cudaMemcpyHostToDevice) != cudaSuccess) {
        set_error("b200_chain_create: descriptor upload failed"); cudaGetLastError();
        cudaFree(ch->d_phases); cudaFree(ch->d_ctl); cudaFree(ch->d_shadow); if (ch->d_ranges) cudaFree(ch->d_ranges); delete ch; return nullptr;
    }
    memset(&ch->p, 0, sizeof(ch->p));
    if (pg) {
        ch->p.total_ctas = pg->total_ctas; ch->p.sync = pg->sync;
        ch->p.n_final = n_final;
        for (int k = 0; k < 4; ++k) ch->p.final_phase[k] = final_phase[k];
        for (int r = 0; r < pg->n_ranks; ++r)
            if (r != pg->rank) ch->p.peer_delta[ch->p.n_peers++] = pg->delta[r];
    }
    ch->p.epoch = ch->d_ctl; ch->p.exit_count = ch->d_ctl + 1;
    ch->p.phases = ch->d_phases; ch->p.n_phases = n_phases; ch->p.ranges = ch->d_ranges;
    ch->p.act_qs_bytes = cfg.act_qs; ch->p.stage_stride = cfg.stage_payload + GS_STAGE_SLACK; ch->p.n_stages = cfg.n_stages;
    ch->p.attn_floats = cfg.attn_floats; ch->p.flags = flags;
    if (flags & 4) {
        // [n_phases][2][8] stamps of CTA 0 / the last CTA, then [n_phases][grid][2] (phase entered, rows consumed) of every CTA
        const size_t n64 = (16 + 2 * (size_t) grid) * (size_t) n_phases;
        if (cudaMalloc((void **) &ch->d_dbg, sizeof(unsigned long long) * n64) == cudaSuccess)
            cudaMemset(ch->d_dbg, 0, sizeof(unsigned long long) * n64);
        ch->p.dbg = ch->d_dbg;
    }
    ch->smem = cfg.smem; ch->grid = grid; ch->t0 = types[0]; ch->t1 = types[1];
    return ch;
}

extern "C" b200_chain * b200_chain_create(const b200_phase * phases, int n_phases, int flags) {
    return gs_chain_create(phases, n_phases, flags, nullptr);
}
extern "C" b200_chain * b200_chain_create_peer(const b200_phase * phases, int n_phases, int flags, const b200_peer_group * pg) {
    if (!pg) { set_error("b200_chain_create_peer: no peer group"); return nullptr; }
    return gs_chain_create(phases, n_phases, flags, pg);
}
extern "C" int b200_chain_grid(void) { return gs_sm_count(); }

// ---------------------------------------------------------------- symmetric arena (CUDA IPC)
extern "C" int b200_peer_alloc(size_t bytes, void ** ptr) {
    if (!ptr || !bytes) return B200_ERR_ARG;
    B200_CUDA_TRY(cudaMalloc(ptr, bytes));
    B200_CUDA_TRY(cudaMemset(*ptr, 0, bytes));
    B200_CUDA_TRY(cudaDeviceSynchronize());
    return B200_OK;
}
extern "C" int b200_peer_free(void * ptr) { B200_CUDA_TRY(cudaFree(ptr)); return B200_OK; }
extern "C" int b200_peer_export(void * ptr, unsigned char handle[64]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size is part of the ABI");
    cudaIpcMemHandle_t h;
    B200_CUDA_TRY(cudaIpcGetMemHandle(&h, ptr));
    memcpy(handle, &h, 64);
    return B200_OK;
}
extern "C" int b200_peer_open(const unsigned char handle[64], void ** ptr) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    B200_CUDA_TRY(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return B200_OK;
}
extern "C" int b200_peer_close(void * ptr) { B200_CUDA_TRY(cudaIpcCloseMemHandle(ptr)); return B200_OK; }

extern "C" int b200_chain_launch(b200_chain * ch, void * stream) {
    if (!ch) return B200_ERR_ARG;
    cudaStream_t st = (cudaStream_t) stream;
    if (ch->peer) {
        // the arrivals of the other ranks for THIS launch cannot come before they have consumed rows this launch produces,
        // and those of the previous launch were all counted before it ended: resetting here does not race with them
        B200_CUDA_TRY(cudaMemsetAsync(ch->p.sync, 0, sizeof(unsigned int) * (size_t) ch->p.n_phases, st));
        return gs_launch_types<2>(ch->p, ch->smem, ch->grid, st, ch->t0, ch->t1);
    }
    return gs_launch_types<1>(ch->p, ch->smem, ch->grid, st, ch->t0, ch->t1);
}

// debug (flags bit 2): copies the globaltimer stamps [phase][cta 0 | last cta][barrier passed, prologue done, consumed, published]
extern "C" int b200_chain_debug_read(b200_chain * ch, unsigned long long * out, int n_phases) {
    if (!ch || !ch->d_dbg || n_phases > ch->p.n_phases) return B200_ERR_ARG;
    B200_CUDA_TRY(cudaMemcpy(out, ch->d_dbg, sizeof(unsigned long long) * 16 * (size_t) n_phases, cudaMemcpyDeviceToHost));
    return B200_OK;
}

// debug (flags bit 2): %globaltimer of every CTA at [phase][cta][phase entered | rows consumed]
extern "C" int b200_chain_debug_read_ctas(b200_chain * ch, unsigned long long * out, int n_phases) {
    if (!ch || !ch->d_dbg || n_phases > ch->p.n_phases) return B200_ERR_ARG;
    B200_CUDA_TRY(cudaMemcpy(out, ch->d_dbg + 16 * (size_t) ch->p.n_phases, sizeof(unsigned long long) * 2 * (size_t) ch->grid * (size_t) n_phases, cudaMemcpyDeviceToHost));
    return B200_OK;
}

extern "C" void b200_chain_destroy(b200_chain * ch) {
    if (!ch) return;
    cudaFree(ch->d_phases); cudaFree(ch->d_ctl); cudaFree(ch->d_shadow); cudaFree(ch->d_ranges); if (ch->d_dbg) cudaFree(ch->d_dbg);
    delete ch;
}
