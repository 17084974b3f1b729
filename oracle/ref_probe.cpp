// Oracle probe translation unit (TEST INFRASTRUCTURE — never linked into the product).
//
// Compiles the reference engine *in place* by including its source file from
// /root/reference (nothing is copied), and adds a few read-only accessors for
// internal stage outputs that the public whisper.h ABI does not expose, so the
// parity tests can compare stage by stage:
//   mel           -> whisper_state::mel            (sys/whisper.cpp/src/whisper.cpp:896)
//   conv output   -> whisper_state::embd_conv      (whisper.cpp:911)
//   encoder out   -> whisper_state::embd_enc       (whisper.cpp:912)
//   cross K/V     -> whisper_state::kv_cross       (whisper.cpp:891)
//   self K/V      -> whisper_state::kv_self        (whisper.cpp:887)
#include "whisper.cpp"

extern "C" {

const float * ref_probe_mel(struct whisper_state * st, int * n_len, int * n_len_org, int * n_mel) {
    *n_len = st->mel.n_len; *n_len_org = st->mel.n_len_org; *n_mel = st->mel.n_mel;
    return st->mel.data.data();
}

static int64_t probe_copy(struct ggml_tensor * t, void * dst, int64_t cap_bytes) {
    if (!t) return -1;
    const int64_t nb = (int64_t) ggml_nbytes(t);
    if (dst && cap_bytes >= nb) ggml_backend_tensor_get(t, dst, 0, nb);
    return nb;
}

// each returns the tensor's byte size (or -1); copies when dst has room. ne[4] receives the ggml shape.
int64_t ref_probe_embd_enc(struct whisper_state * st, void * dst, int64_t cap, int64_t * ne) {
    if (st->embd_enc && ne) for (int i = 0; i < 4; ++i) ne[i] = st->embd_enc->ne[i];
    return probe_copy(st->embd_enc, dst, cap);
}
int64_t ref_probe_embd_conv(struct whisper_state * st, void * dst, int64_t cap, int64_t * ne) {
    if (st->embd_conv && ne) for (int i = 0; i < 4; ++i) ne[i] = st->embd_conv->ne[i];
    return probe_copy(st->embd_conv, dst, cap);
}
int64_t ref_probe_kv_cross_k(struct whisper_state * st, void * dst, int64_t cap) { return probe_copy(st->kv_cross.k, dst, cap); }
int64_t ref_probe_kv_cross_v(struct whisper_state * st, void * dst, int64_t cap) { return probe_copy(st->kv_cross.v, dst, cap); }
int64_t ref_probe_kv_self_k (struct whisper_state * st, void * dst, int64_t cap) { return probe_copy(st->kv_self.k,  dst, cap); }
int64_t ref_probe_kv_self_v (struct whisper_state * st, void * dst, int64_t cap) { return probe_copy(st->kv_self.v,  dst, cap); }
int     ref_probe_kv_self_size(struct whisper_state * st) { return (int) st->kv_self.size; }

// counters the reference keeps per state (whisper.cpp:868-881)
void ref_probe_counters(struct whisper_state * st, int32_t * out7, int64_t * us6) {
    out7[0] = st->n_sample; out7[1] = st->n_encode; out7[2] = st->n_decode; out7[3] = st->n_batchd;
    out7[4] = st->n_prompt; out7[5] = st->n_fail_p; out7[6] = st->n_fail_h;
    us6[0] = st->t_sample_us; us6[1] = st->t_encode_us; us6[2] = st->t_decode_us;
    us6[3] = st->t_batchd_us; us6[4] = st->t_prompt_us; us6[5] = st->t_mel_us;
}

// last decode's alignment-head cross-attention weights (whisper.cpp:947-948)
const float * ref_probe_aheads_qks(struct whisper_state * st, int64_t * n) {
    *n = (int64_t) st->aheads_cross_QKs_data.size();
    return st->aheads_cross_QKs_data.data();
}

}
