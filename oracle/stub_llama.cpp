// TEST INFRASTRUCTURE ONLY — a scripted stand-in for libllama.
//
// north_star puts llama.cpp token generation out of scope, but the reference's streaming policy (emit_range /
// maybe_emit, /root/reference/src/test-to-speech.cpp:367-417, :499-571, :574-618) lives INSIDE the LLM branch of
// TestToSpeech::synthesize_stream_profiled and cannot run without the llama_* symbols.  This file provides exactly the
// 23 symbols src/test-to-speech.cpp references, with llama.h's own signatures (it is compiled against the reference's
// header), so that the UNMODIFIED src/test-to-speech.cpp links and runs its LLM branch on the CPU:
//
//   * the "model file" is a text file of integers: the speech codes the "LLM" will emit, one per generated token
//   * llama_sampler_sample returns the next scripted code as a token, then EOS
//   * llama_token_to_piece renders token N as the piece "<|s_N|>", which the reference's own parse_speech_tokens
//     turns back into code N
//
// Nothing of llama.cpp's implementation is used or copied; the prompt / tokenizer / KV-cache calls are accepted and
// ignored.  Used by oracle/_ref/libmio_ref_*.so (the streaming oracle) and by the drop-in binaries of
// oracle/build_dropin.sh (the reference's executables linked against our shim).
#include "llama.h"   // the reference's header: /root/reference/third_party/llama.cpp/include

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

struct llama_vocab { int unused = 0; };
struct llama_model {
    std::vector<int> script;
    llama_vocab vocab;
};
struct llama_context {
    const llama_model * model = nullptr;
    size_t next = 0;
};

static const llama_token STUB_EOS = 1 << 24;
static const llama_token STUB_IM_END = (1 << 24) + 1;

extern "C" {

void llama_backend_init(void) {}
void llama_backend_free(void) {}

struct llama_model_params llama_model_default_params(void) {
    struct llama_model_params p;
    memset(&p, 0, sizeof(p));
    return p;
}
struct llama_context_params llama_context_default_params(void) {
    struct llama_context_params p;
    memset(&p, 0, sizeof(p));
    return p;
}
struct llama_sampler_chain_params llama_sampler_chain_default_params(void) {
    struct llama_sampler_chain_params p;
    memset(&p, 0, sizeof(p));
    return p;
}

struct llama_model * llama_model_load_from_file(const char * path_model, struct llama_model_params) {
    FILE * f = fopen(path_model, "r");
    if (!f) return nullptr;
    llama_model * m = new llama_model();
    int v;
    while (fscanf(f, "%d", &v) == 1) m->script.push_back(v);
    fclose(f);
    return m;
}
void llama_model_free(struct llama_model * model) { delete model; }
const struct llama_vocab * llama_model_get_vocab(const struct llama_model * model) { return &model->vocab; }

struct llama_context * llama_init_from_model(struct llama_model * model, struct llama_context_params) {
    llama_context * c = new llama_context();
    c->model = model;
    return c;
}
void llama_free(struct llama_context * ctx) { delete ctx; }

// one token per prompt byte (capped); "<|im_end|>" alone is the single special token the reference looks for
int32_t llama_tokenize(const struct llama_vocab *, const char * text, int32_t text_len, llama_token * tokens,
                       int32_t n_tokens_max, bool, bool) {
    if (text_len == 10 && strncmp(text, "<|im_end|>", 10) == 0) {
        if (n_tokens_max < 1) return -1;
        tokens[0] = STUB_IM_END;
        return 1;
    }
    const int32_t n = text_len < n_tokens_max ? text_len : n_tokens_max;
    for (int32_t i = 0; i < n; i++) tokens[i] = (llama_token) (unsigned char) text[i];
    return n;
}

struct llama_sampler * llama_sampler_chain_init(struct llama_sampler_chain_params) {
    llama_sampler * s = new llama_sampler();
    s->iface = nullptr;
    s->ctx = nullptr;
    return s;
}
void llama_sampler_chain_add(struct llama_sampler *, struct llama_sampler * smpl) { delete smpl; }
struct llama_sampler * llama_sampler_init_temp(float) { return llama_sampler_chain_init(llama_sampler_chain_default_params()); }
struct llama_sampler * llama_sampler_init_dist(uint32_t) { return llama_sampler_chain_init(llama_sampler_chain_default_params()); }
void llama_sampler_free(struct llama_sampler * smpl) { delete smpl; }
void llama_sampler_accept(struct llama_sampler *, llama_token) {}

llama_token llama_sampler_sample(struct llama_sampler *, struct llama_context * ctx, int32_t) {
    if (ctx->next >= ctx->model->script.size()) return STUB_EOS;
    return (llama_token) ctx->model->script[ctx->next++];
}

struct llama_batch llama_batch_init(int32_t n_tokens, int32_t, int32_t n_seq_max) {
    llama_batch b;
    memset(&b, 0, sizeof(b));
    const int n = n_tokens > 0 ? n_tokens : 1;
    b.token = (llama_token *) calloc(n, sizeof(llama_token));
    b.pos = (llama_pos *) calloc(n, sizeof(llama_pos));
    b.n_seq_id = (int32_t *) calloc(n, sizeof(int32_t));
    b.seq_id = (llama_seq_id **) calloc(n + 1, sizeof(llama_seq_id *));
    for (int i = 0; i < n; i++) b.seq_id[i] = (llama_seq_id *) calloc(n_seq_max > 0 ? n_seq_max : 1, sizeof(llama_seq_id));
    b.logits = (int8_t *) calloc(n, sizeof(int8_t));
    return b;
}
void llama_batch_free(struct llama_batch b) {
    free(b.token); free(b.pos); free(b.n_seq_id); free(b.logits);
    if (b.seq_id) { for (int i = 0; b.seq_id[i]; i++) free(b.seq_id[i]); free(b.seq_id); }
}
int32_t llama_decode(struct llama_context *, struct llama_batch) { return 0; }

llama_token llama_vocab_eos(const struct llama_vocab *) { return STUB_EOS; }

int32_t llama_token_to_piece(const struct llama_vocab *, llama_token token, char * buf, int32_t length, int32_t, bool) {
    char tmp[64];
    const int n = snprintf(tmp, sizeof(tmp), "<|s_%d|>", (int) token);
    if (n > length) return -n;
    memcpy(buf, tmp, n);
    return n;
}

}  // extern "C"
