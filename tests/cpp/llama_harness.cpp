// llama_harness.cpp -- TEST INFRASTRUCTURE.  Drives the REFERENCE's own, unmodified llama.cpp (oracle/_ref/libllama_ref.so:
// llama_load_model_from_file -> llama_new_context_with_model -> llama_decode, src/llama.cpp:21207 / :17114, under
// ggml_backend_sched, ggml-backend.cpp:2108-2329) over a GGUF file, greedy-decodes pp + tg the way koboldcpp's benchmark does
// (koboldcpp.py:4276-4345) and prints one JSON line: token stream, top-1/top-2 margins, tok/s.
//
//   --backend-lib <libkcpp_b200.so>   dlopen the B200 backend and attach it at run time (Route 2):
//                                     ggml_backend_register(ggml_backend_b200_reg())  (ggml-backend-impl.h:220)
//   without it the run is the reference CPU backend alone = the oracle / CPU baseline of the llama_decode level.
//
// Built by oracle/Makefile against the reference headers where they lie (nothing is copied), output in oracle/_ref/.
#include <dlfcn.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "ggml-backend-impl.h"
#include "ggml-backend.h"
#include "ggml.h"
#include "llama.h"

static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

struct Args {
    std::string model, backend_lib, logits_out, state_save, state_load, split_mode = "none", type_kv = "f16";
    int ngl = 999, n_prompt = 128, n_gen = 128, n_ctx = 0, threads = 0, fa = 1, n_batch = 512, n_ubatch = 512, main_gpu = 0, shift_at = -1, shift_by = 0, n_warm = 0, kv_fill = 0;
    bool quiet = false;
};

static void log_cb(enum ggml_log_level level, const char *text, void *ud) {
    const bool quiet = *(const bool *)ud;
    if (!quiet || level == GGML_LOG_LEVEL_ERROR) fputs(text, stderr);
}

int main(int argc, char **argv) {
    Args a;
    for (int i = 1; i < argc; i++) {
        std::string k = argv[i];
        auto val = [&]() -> const char * { if (i + 1 >= argc) { fprintf(stderr, "missing value for %s\n", k.c_str()); exit(2); } return argv[++i]; };
        if (k == "--model") a.model = val();
        else if (k == "--backend-lib") a.backend_lib = val();
        else if (k == "--ngl") a.ngl = atoi(val());
        else if (k == "--n-prompt") a.n_prompt = atoi(val());
        else if (k == "--n-gen") a.n_gen = atoi(val());
        else if (k == "--n-ctx") a.n_ctx = atoi(val());
        else if (k == "--threads") a.threads = atoi(val());
        else if (k == "--fa") a.fa = atoi(val());
        else if (k == "--n-batch") a.n_batch = atoi(val());
        else if (k == "--n-ubatch") a.n_ubatch = atoi(val());
        else if (k == "--type-kv") a.type_kv = val();
        else if (k == "--split-mode") a.split_mode = val();
        else if (k == "--main-gpu") a.main_gpu = atoi(val());
        else if (k == "--logits-out") a.logits_out = val();
        else if (k == "--state-save") a.state_save = val();
        else if (k == "--state-load") a.state_load = val();
        else if (k == "--shift-at") a.shift_at = atoi(val());   // after this many generated tokens: drop `shift-by` oldest cells + K-shift
        else if (k == "--shift-by") a.shift_by = atoi(val());
        else if (k == "--n-warm") a.n_warm = atoi(val());      // untimed decode steps in front of the n_gen timed ones
        else if (k == "--kv-fill") a.kv_fill = atoi(val());    // after the prompt: replicate its KV rows so that kv_fill cells are occupied
        else if (k == "--quiet") a.quiet = true;
        else { fprintf(stderr, "unknown argument %s\n", k.c_str()); return 2; }
    }
    if (a.model.empty()) { fprintf(stderr, "--model required\n"); return 2; }
    llama_log_set(log_cb, &a.quiet);

    int n_dev = 0;
    if (!a.backend_lib.empty()) {
        void *h = dlopen(a.backend_lib.c_str(), RTLD_NOW | RTLD_LOCAL);
        if (!h) { fprintf(stderr, "dlopen %s: %s\n", a.backend_lib.c_str(), dlerror()); return 3; }
        typedef ggml_backend_reg_t (*reg_fn)(void);
        reg_fn reg = (reg_fn)dlsym(h, "ggml_backend_b200_reg");
        if (!reg) { fprintf(stderr, "ggml_backend_b200_reg not exported\n"); return 3; }
        ggml_backend_reg_t r = reg();
        n_dev = (int)r->iface.get_device_count(r);
        if (n_dev <= 0) { fprintf(stderr, "the B200 backend reports no device (no CPU fallback)\n"); return 4; }
        ggml_backend_register(r);
    }
    llama_backend_init();

    llama_model_params mp = llama_model_default_params();
    mp.n_gpu_layers = a.backend_lib.empty() ? 0 : a.ngl;
    mp.split_mode = a.split_mode == "row" ? LLAMA_SPLIT_MODE_ROW : a.split_mode == "layer" ? LLAMA_SPLIT_MODE_LAYER : LLAMA_SPLIT_MODE_NONE;
    mp.main_gpu = a.main_gpu;
    mp.use_mmap = true;
    const double t_load0 = now_s();
    llama_model *model = llama_load_model_from_file(a.model.c_str(), mp);
    if (!model) { fprintf(stderr, "failed to load %s\n", a.model.c_str()); return 5; }
    const double t_load = now_s() - t_load0;

    llama_context_params cp = llama_context_default_params();
    cp.n_ctx = a.n_ctx > 0 ? a.n_ctx : (uint32_t)(std::max(a.n_prompt, a.kv_fill) + a.n_gen + a.n_warm + 8);
    cp.n_batch = a.n_batch;
    cp.n_ubatch = a.n_ubatch;
    if (a.threads > 0) { cp.n_threads = a.threads; cp.n_threads_batch = a.threads; }
    cp.flash_attn = a.fa != 0;
    cp.type_k = cp.type_v = a.type_kv == "q8_0" ? GGML_TYPE_Q8_0 : a.type_kv == "q4_0" ? GGML_TYPE_Q4_0 : GGML_TYPE_F16;
    cp.offload_kqv = true;
    cp.no_perf = true;
    llama_context *ctx = llama_new_context_with_model(model, cp);
    if (!ctx) { fprintf(stderr, "failed to create the context\n"); return 6; }

    const int n_vocab = llama_n_vocab(model);
    std::vector<llama_token> prompt((size_t)a.n_prompt);
    for (int i = 0; i < a.n_prompt; i++) prompt[(size_t)i] = (llama_token)(((int64_t)i * 7919 + 1) % n_vocab);  // fixed ids

    auto argmax2 = [&](const float *lg, float *margin) {
        int best = 0;
        float b1 = lg[0], b2 = -INFINITY;
        for (int i = 1; i < n_vocab; i++) {
            if (lg[i] > b1) { b2 = b1; b1 = lg[i]; best = i; }
            else if (lg[i] > b2) b2 = lg[i];
        }
        *margin = b1 - b2;
        return best;
    };
    FILE *lf = a.logits_out.empty() ? nullptr : fopen(a.logits_out.c_str(), "wb");

    int n_past = 0;
    llama_token tok = 0;
    float margin = 0.f;
    std::vector<int> tokens;
    std::vector<float> margins, lstd;
    double t_pp = 0.0;
    auto take_logits = [&]() {
        const float *lg = llama_get_logits_ith(ctx, -1);
        if (lf) fwrite(lg, sizeof(float), (size_t)n_vocab, lf);
        double s = 0, ss = 0;
        for (int i = 0; i < n_vocab; i++) { s += lg[i]; ss += (double)lg[i] * lg[i]; }
        lstd.push_back((float)std::sqrt(std::max(0.0, ss / n_vocab - (s / n_vocab) * (s / n_vocab))));
        tok = argmax2(lg, &margin);
        tokens.push_back(tok);
        margins.push_back(margin);
    };

    if (!a.state_load.empty()) {
        FILE *f = fopen(a.state_load.c_str(), "rb");
        if (!f) { fprintf(stderr, "cannot open %s\n", a.state_load.c_str()); return 7; }
        int32_t hdr[2];
        uint64_t sz = 0;
        if (fread(hdr, 4, 2, f) != 2 || fread(&sz, 8, 1, f) != 1) return 7;
        std::vector<uint8_t> buf((size_t)sz);
        if (fread(buf.data(), 1, (size_t)sz, f) != (size_t)sz) return 7;
        fclose(f);
        if (llama_state_set_data(ctx, buf.data(), buf.size()) == 0) { fprintf(stderr, "llama_state_set_data failed\n"); return 7; }
        tok = hdr[0];
        n_past = hdr[1];
        tokens.push_back(tok);
        margins.push_back(0.f);
        lstd.push_back(0.f);
    } else {
        const double t0 = now_s();
        for (int i0 = 0; i0 < a.n_prompt; i0 += a.n_batch) {
            const int n = std::min(a.n_batch, a.n_prompt - i0);
            if (llama_decode(ctx, llama_batch_get_one(prompt.data() + i0, n, n_past, 0)) != 0) { fprintf(stderr, "llama_decode (prompt) failed\n"); return 8; }
            n_past += n;
        }
        llama_synchronize(ctx);
        t_pp = now_s() - t0;
        take_logits();
        if (!a.state_save.empty()) {
            std::vector<uint8_t> buf(llama_state_get_size(ctx));
            const size_t n = llama_state_get_data(ctx, buf.data(), buf.size());
            FILE *f = fopen(a.state_save.c_str(), "wb");
            int32_t hdr[2] = {tok, n_past};
            uint64_t sz = n;
            fwrite(hdr, 4, 2, f); fwrite(&sz, 8, 1, f); fwrite(buf.data(), 1, n, f);
            fclose(f);
        }
    }

    if (a.kv_fill > n_past) {
        // "KV pre-filled to N cells"  without computing an N-token prompt on the CPU: take the session state
        // the reference itself serialises (llama_state_get_data, src/llama.cpp:20825-20845: arch, output ids, logits, embeddings, then
        // cell_count, per-cell {pos, n_seq_id, seq_ids}, v_trans, n_layer, per layer {type, row size, rows} for K and for V), repeat
        // the prompt's cell rows cyclically up to N cells with positions 0..N-1, and hand it back through llama_state_set_data.
        std::vector<uint8_t> st(llama_state_get_size(ctx));
        const size_t got = llama_state_get_data(ctx, st.data(), st.size());
        size_t off = 0;
        auto rd32 = [&](size_t o) { uint32_t v; memcpy(&v, st.data() + o, 4); return v; };
        auto rd64 = [&](size_t o) { uint64_t v; memcpy(&v, st.data() + o, 8); return v; };
        off += 4 + rd32(off);                         // arch string
        off += 4 + 4 * (size_t)rd32(off);             // output ids
        off += 8 + 4 * (size_t)rd64(off);             // logits
        off += 8 + 4 * (size_t)rd64(off);             // embeddings
        const size_t hdr_end = off;
        const uint32_t n0 = rd32(off); off += 4;
        if ((int)n0 != n_past || got < off) { fprintf(stderr, "--kv-fill: unexpected state layout (cells %u, n_past %d)\n", n0, n_past); return 10; }
        for (uint32_t i = 0; i < n0; i++) { const uint32_t ns = rd32(off + 4); off += 8 + 4 * (size_t)ns; }
        const uint32_t v_trans = rd32(off), n_layer = rd32(off + 4); off += 8;
        if (v_trans) { fprintf(stderr, "--kv-fill needs flash attention (non-transposed V cache)\n"); return 10; }
        std::vector<uint8_t> out(st.begin(), st.begin() + hdr_end);
        auto put = [&](const void *p, size_t n) { const uint8_t *b = (const uint8_t *)p; out.insert(out.end(), b, b + n); };
        const uint32_t N = (uint32_t)a.kv_fill;
        put(&N, 4);
        for (uint32_t i = 0; i < N; i++) { const int32_t pos = (int32_t)i; const uint32_t ns = 1; const int32_t sid = 0; put(&pos, 4); put(&ns, 4); put(&sid, 4); }
        put(&v_trans, 4); put(&n_layer, 4);
        for (int kv = 0; kv < 2; kv++) {
            for (uint32_t il = 0; il < n_layer; il++) {
                const uint64_t row = rd64(off + 4);
                put(st.data() + off, 12);
                off += 12;
                for (uint32_t i = 0; i < N; i++) put(st.data() + off + (size_t)(i % n0) * row, row);
                off += (size_t)n0 * row;
            }
        }
        if (llama_state_set_data(ctx, out.data(), out.size()) == 0) { fprintf(stderr, "--kv-fill: llama_state_set_data failed\n"); return 10; }
        n_past = a.kv_fill;
    }

    // tg: n_gen single-token decodes (the first generated token came from the prompt logits); n_warm untimed ones in front
    double t1 = now_s();
    int n_dec = 0;
    for (int g = -a.n_warm; g < a.n_gen; g++) {
        if (g == 0) { llama_synchronize(ctx); t1 = now_s(); n_dec = 0; }
        if (a.shift_at >= 0 && g == a.shift_at && a.shift_by > 0) {
            // koboldcpp ContextShift: drop the oldest cells, slide the rest (llama_kv_cache_seq_rm + seq_add -> K-shift graph,
            // src/llama.cpp:16530 build_k_shift) -- gpttype_adapter.cpp PurgeMissingTokens / context shifting
            llama_kv_cache_seq_rm(ctx, 0, 0, a.shift_by);
            llama_kv_cache_seq_add(ctx, 0, a.shift_by, -1, -a.shift_by);
            n_past -= a.shift_by;
        }
        llama_token t = tok;
        if (llama_decode(ctx, llama_batch_get_one(&t, 1, n_past, 0)) != 0) { fprintf(stderr, "llama_decode (gen %d) failed\n", g); return 9; }
        n_past += 1;
        n_dec++;
        take_logits();
    }
    llama_synchronize(ctx);
    const double t_tg = now_s() - t1;
    if (lf) fclose(lf);

    printf("{\"backend\": \"%s\", \"n_devices\": %d, \"n_vocab\": %d, \"n_prompt\": %d, \"n_gen\": %d, \"n_ctx\": %u, \"threads\": %d, \"load_s\": %.3f, "
           "\"pp_s\": %.6f, \"pp_tok_s\": %.3f, \"tg_s\": %.6f, \"tg_tok_s\": %.3f, \"tokens\": [",
           a.backend_lib.empty() ? "cpu" : "b200", n_dev, n_vocab, a.n_prompt, n_dec, llama_n_ctx(ctx), llama_n_threads(ctx), t_load, t_pp,
           t_pp > 0 ? a.n_prompt / t_pp : 0.0, t_tg, t_tg > 0 ? n_dec / t_tg : 0.0);
    for (size_t i = 0; i < tokens.size(); i++) printf("%s%d", i ? ", " : "", tokens[i]);
    printf("], \"margins\": [");
    for (size_t i = 0; i < margins.size(); i++) printf("%s%.5g", i ? ", " : "", margins[i]);
    printf("], \"logit_std\": [");
    for (size_t i = 0; i < lstd.size(); i++) printf("%s%.5g", i ? ", " : "", lstd[i]);
    printf("]}\n");
    fflush(stdout);

    llama_free(ctx);
    llama_free_model(model);
    llama_backend_free();
    return 0;
}
