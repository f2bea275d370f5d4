// oracle/plugin_parity.cpp — TEST INFRASTRUCTURE.  End-to-end parity of the B200 plugin against the reference's CPU
// backend THROUGH the reference's own public API (llama.h): the same GGUF model decodes the same tokens once on the CPU
// backend and once with every layer offloaded to the plugin; the logits of every step are compared (NMSE).
// usage: plugin-parity model.gguf backend_dir plugin.so [n_prompt n_decode tol]
// B200_PARITY_SPLIT_ROW=1: the plugin run uses --split-mode row (weights in the backend's split buffer type).
// B200_PARITY_KV=q8_0 (or f16): type of the K and V caches in BOTH runs (-ctk / -ctv); B200_PARITY_NCTX=n: context size
// (default 1024).  Prompts longer than the batch size are decoded in chunks; the logits of the last chunk are compared.
#include "llama.h"
#include "ggml.h"
#include "ggml-backend.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <cstdio>
#include <cstdlib>
#include <vector>

static std::vector<std::vector<float>> run(const char * path, int ngl, int n_prompt, int n_decode) {
    llama_model_params mp = llama_model_default_params();
    mp.n_gpu_layers = ngl < 0 ? 0 : ngl;
    if (ngl > 0 && getenv("B200_PARITY_SPLIT_ROW")) mp.split_mode = LLAMA_SPLIT_MODE_ROW;
    llama_model * m = llama_model_load_from_file(path, mp);
    if (!m) { fprintf(stderr, "model load failed\n"); exit(3); }
    const int n_vocab = llama_vocab_n_tokens(llama_model_get_vocab(m));
    llama_context_params cp = llama_context_default_params();
    cp.n_ctx = getenv("B200_PARITY_NCTX") ? atoi(getenv("B200_PARITY_NCTX")) : 1024;
    cp.n_batch = 512; cp.flash_attn_type = LLAMA_FLASH_ATTN_TYPE_ENABLED; cp.n_threads = 8; cp.n_threads_batch = 8;
    const char * kv_override = ngl == 0 ? nullptr : getenv("B200_PARITY_KV2");     // second run only (sensitivity probe, below)
    if (const char * kv = kv_override ? kv_override : getenv("B200_PARITY_KV")) {
        const ggml_type t = !strcmp(kv, "q8_0") ? GGML_TYPE_Q8_0 : !strcmp(kv, "q4_0") ? GGML_TYPE_Q4_0 : GGML_TYPE_F16;
        cp.type_k = t; cp.type_v = t;
    }
    llama_context * c = llama_init_from_model(m, cp);
    if (!c) { fprintf(stderr, "context creation failed\n"); exit(3); }
    std::vector<std::vector<float>> out;
    std::vector<llama_token> prompt;
    for (int i = 0; i < n_prompt; ++i) prompt.push_back((llama_token) ((i * 37 + 11) % n_vocab));
    for (int i0 = 0; i0 < n_prompt; i0 += (int) cp.n_batch)
        if (llama_decode(c, llama_batch_get_one(prompt.data() + i0, std::min<int>(cp.n_batch, n_prompt - i0))) != 0) { fprintf(stderr, "prompt decode failed\n"); exit(4); }
    { const float * l = llama_get_logits_ith(c, -1); out.emplace_back(l, l + n_vocab); }
    for (int i = 0; i < n_decode; ++i) {
        llama_token t = (llama_token) ((i * 101 + 7) % n_vocab);     // fixed stream: both runs see identical inputs
        if (llama_decode(c, llama_batch_get_one(&t, 1)) != 0) { fprintf(stderr, "decode failed\n"); exit(4); }
        const float * l = llama_get_logits_ith(c, -1);
        out.emplace_back(l, l + n_vocab);
    }
    llama_free(c);
    llama_model_free(m);
    return out;
}

int main(int argc, char ** argv) {
    if (argc < 4) { fprintf(stderr, "usage: plugin-parity model.gguf backend_dir plugin.so [n_prompt n_decode tol]\n"); return 2; }
    const int n_prompt = argc > 4 ? atoi(argv[4]) : 5, n_decode = argc > 5 ? atoi(argv[5]) : 12;
    const double tol = argc > 6 ? atof(argv[6]) : 2e-3;
    ggml_backend_load_all_from_path(argv[2]);
    const auto ref = run(argv[1], 0, n_prompt, n_decode);            // only CPU devices are registered at this point
    // plugin.so == "cpu": sensitivity probe of the MODEL, not a parity run: the second run is the CPU backend again, with
    // the KV-cache type of B200_PARITY_KV2 (e.g. q8_0: a relative perturbation of ~0.4 % on K and V).  The NMSE it prints
    // is how much the logits of this model move under a perturbation of the size of one quantisation step.
    const bool probe = !strcmp(argv[3], "cpu");
    if (!probe && !ggml_backend_load(argv[3])) { fprintf(stderr, "cannot load plugin %s\n", argv[3]); return 5; }
    const auto got = run(argv[1], probe ? -1 : 999, n_prompt, n_decode);
    int bad = 0;
    for (size_t s = 0; s < ref.size(); ++s) {
        double num = 0, den = 0; int am_r = 0, am_g = 0;
        for (size_t i = 0; i < ref[s].size(); ++i) {
            const double d = (double) got[s][i] - ref[s][i];
            num += d * d; den += (double) ref[s][i] * ref[s][i];
            if (ref[s][i] > ref[s][am_r]) am_r = (int) i;
            if (got[s][i] > got[s][am_g]) am_g = (int) i;
        }
        const double nmse = num / (den > 0 ? den : 1);
        const bool ok = std::isfinite(nmse) && nmse <= tol;
        bad += !ok;
        printf("step %2zu %-7s nmse %.3e argmax ref %d got %d %s\n", s, s == 0 ? "prompt" : "decode", nmse, am_r, am_g, ok ? "ok" : "FAIL");
    }
    printf("plugin-parity: %s (%d of %zu steps over tolerance %.1e)\n", bad ? "FAIL" : "PASS", bad, ref.size(), tol);
    return bad ? 1 : 0;
}
