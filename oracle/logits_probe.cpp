// logits_probe.cpp -- TEST INFRASTRUCTURE (oracle/): a few lines against the reference's public llama.h API, linked with the reference's own
// libllama / libggml compiled in place (oracle/_ref/avx2, `make -C oracle llama`).  It evaluates one fixed token sequence with the
// reference runtime -- a prompt batch, then single-token decode steps that cross a KV-cache view boundary -- and writes the logits of
// every step to a file.  tests/test_gpu_llama_logits.py runs it twice on the same random-block GGUF: on ggml-cpu (-ngl 0) and through
// the B200 plugin (GGML_BACKEND_PATH, -ngl 99), and compares: the end-to-end parity check of the drop-in, through the reference's
// graph builder, scheduler, allocator and KV cache (include/llama.h:437-996 for the calls used).
// usage: logits_probe <model.gguf> <out.bin> <ngl> <n_prompt> <n_decode> [tokens per decode step = 1]
#include <string.h>
#include "llama.h"
#include "ggml-backend.h"
#include <cstdio>
#include <cstdlib>
#include <vector>

int main(int argc, char ** argv) {
    if (argc < 6) { fprintf(stderr, "usage: %s model out ngl n_prompt n_decode\n", argv[0]); return 2; }
    const int ngl = atoi(argv[3]), n_prompt = atoi(argv[4]), n_decode = atoi(argv[5]), n_step = argc > 6 ? atoi(argv[6]) : 1;
    ggml_backend_load_all();
    llama_backend_init();
    llama_model_params mp = llama_model_default_params();
    mp.n_gpu_layers = ngl;
    if (getenv("LOGITS_PROBE_SPLIT_ROW")) mp.split_mode = LLAMA_SPLIT_MODE_ROW;   // -sm row: weight matrices in the backend's split buffer type (src/llama-model.cpp:393-411)
    mp.use_mmap = false;                                   // the loader's chunked async upload path: exercises the plugin's load-time repack hook
    llama_model * model = llama_model_load_from_file(argv[1], mp);
    if (!model) { fprintf(stderr, "logits_probe: cannot load %s\n", argv[1]); return 1; }
    llama_context_params cp = llama_context_default_params();
    cp.n_ctx = 2048; cp.n_batch = 1024; cp.n_ubatch = 512;   // prompts up to 1024 tokens, evaluated as ubatches of 512
    cp.n_threads = 8; cp.n_threads_batch = 8;
    cp.flash_attn_type = getenv("LOGITS_PROBE_NO_FA") ? LLAMA_FLASH_ATTN_TYPE_DISABLED : LLAMA_FLASH_ATTN_TYPE_ENABLED;   // -fa 0: attention as MUL_MAT + SOFT_MAX + MUL_MAT, V cache transposed
    cp.no_perf = true;
    auto kv_type = [](const char * e, ggml_type def) {           // -ctk / -ctv: the KV cache as q8_0 / q4_0 blocks (SET_ROWS quantizes, FLASH_ATTN_EXT reads them)
        const char * v = getenv(e);
        if (!v) return def;
        return !strcmp(v, "q8_0") ? GGML_TYPE_Q8_0 : !strcmp(v, "q4_0") ? GGML_TYPE_Q4_0 : def;
    };
    cp.type_k = kv_type("LOGITS_PROBE_CTK", cp.type_k); cp.type_v = kv_type("LOGITS_PROBE_CTV", cp.type_v);
    llama_context * ctx = llama_init_from_model(model, cp);
    if (!ctx) { fprintf(stderr, "logits_probe: cannot create a context\n"); return 1; }
    const int n_vocab = llama_vocab_n_tokens(llama_model_get_vocab(model));
    FILE * out = fopen(argv[2], "wb");
    if (!out) return 1;
    unsigned s = 777;
    auto next_tok = [&]() { s = s * 1664525u + 1013904223u; return (llama_token) ((s >> 8) % (unsigned) n_vocab); };
    std::vector<llama_token> prompt(n_prompt);
    for (auto & t : prompt) t = next_tok();
    if (llama_decode(ctx, llama_batch_get_one(prompt.data(), n_prompt)) != 0) { fprintf(stderr, "logits_probe: prompt decode failed\n"); return 1; }
    fwrite(llama_get_logits_ith(ctx, -1), sizeof(float), n_vocab, out);
    for (int i = 0; i < n_decode; ++i) {
        std::vector<llama_token> t(n_step);                 // n_step > 1: a few positions of the same sequence per step (speculative-style small batch)
        for (auto & x : t) x = next_tok();
        if (llama_decode(ctx, llama_batch_get_one(t.data(), n_step)) != 0) { fprintf(stderr, "logits_probe: decode step %d failed\n", i); return 1; }
        fwrite(llama_get_logits_ith(ctx, -1), sizeof(float), n_vocab, out);
    }
    fclose(out);
    printf("logits_probe: wrote %d x %d logits (ngl %d)\n", 1 + n_decode, n_vocab, ngl);
    llama_free(ctx); llama_model_free(model); llama_backend_free();
    return 0;
}
