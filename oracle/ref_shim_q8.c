// TEST INFRASTRUCTURE ONLY -- never linked into the product path.
//
// Thin C shim around the UNMODIFIED reference Q8_0 CPU path (runq.c), compiled
// where it lies.  `#define TESTING` removes its main() (guard at runq.c:1008).
// Visible afterwards: quantize runq.c:145-171, dequantize :139-143, matmul
// :317-342 (OpenMP over rows), forward :344-481, loader :174-258.
// runq.c is C99 (designated initialisers, implicit void* casts): this TU must
// be compiled as C.  Built by oracle/Makefile into oracle/_ref/libref_q8.so.
#define TESTING 1
#include REF_RUNQ_C   // -DREF_RUNQ_C='"/root/reference/runq.c"'

#include <omp.h>

void* refq8_open(const char* checkpoint_path) {
    Transformer* t = (Transformer*)calloc(1, sizeof(Transformer));
    build_transformer(t, (char*)checkpoint_path);
    return t;
}

void* refq8_clone(void* h) {
    Transformer* src = (Transformer*)h;
    Transformer* t = (Transformer*)calloc(1, sizeof(Transformer));
    t->config = src->config;
    t->weights = src->weights;
    t->fd = -1;
    t->data = (float*)MAP_FAILED;
    malloc_run_state(&t->state, &t->config);
    return t;
}

void refq8_config(void* h, int out[8]) {
    Config* c = &((Transformer*)h)->config;
    out[0] = c->dim; out[1] = c->hidden_dim; out[2] = c->n_layers; out[3] = c->n_heads;
    out[4] = c->n_kv_heads; out[5] = c->vocab_size; out[6] = c->seq_len; out[7] = GS;
}

int refq8_group_size(void) { return GS; }
void refq8_set_group_size(int gs) { GS = gs; }

float* refq8_forward(void* h, int token, int pos) {
    return forward((Transformer*)h, token, pos);
}

// which: 0 x,1 xb,2 xb2,3 hb,4 hb2,5 q,6 att,7 logits,8 key_cache,9 value_cache
float* refq8_state_ptr(void* h, int which) {
    RunState* s = &((Transformer*)h)->state;
    switch (which) {
        case 0: return s->x;   case 1: return s->xb;  case 2: return s->xb2;
        case 3: return s->hb;  case 4: return s->hb2; case 5: return s->q;
        case 6: return s->att; case 7: return s->logits;
        case 8: return s->key_cache; case 9: return s->value_cache;
    }
    return NULL;
}

// Last activation quantised by forward(): which 0 -> xq (dim), 1 -> hq (hidden_dim)
int8_t* refq8_state_q(void* h, int which) {
    RunState* s = &((Transformer*)h)->state;
    return which == 0 ? s->xq.q : s->hq.q;
}
float* refq8_state_s(void* h, int which) {
    RunState* s = &((Transformer*)h)->state;
    return which == 0 ? s->xq.s : s->hq.s;
}

// runq.c:145-171, group size taken from the global GS
void refq8_quantize(int8_t* q, float* s, float* x, int n) {
    QuantizedTensor qt; qt.q = q; qt.s = s;
    quantize(&qt, x, n);
}
void refq8_dequantize(int8_t* q, float* s, float* x, int n) {
    QuantizedTensor qt; qt.q = q; qt.s = s;
    dequantize(&qt, x, n);
}
// runq.c:317-342: xout[d] = W[d,n] (int8 + scales) . x[n] (int8 + scales)
void refq8_matmul(float* xout, int8_t* xq, float* xs, int8_t* wq, float* ws, int n, int d) {
    QuantizedTensor x; x.q = xq; x.s = xs;
    QuantizedTensor w; w.q = wq; w.s = ws;
    matmul(xout, &x, &w, n, d);
}

// n independent sequences, one step each.  runq.c parallelises inside matmul
// (omp parallel for over rows) and over heads; sequences run one after another
// with the team size given here, exactly as the shipped binary would.
void refq8_forward_many(void** hs, const int* token, const int* pos, int n, int n_threads) {
    if (n_threads > 0) omp_set_num_threads(n_threads);
    for (int i = 0; i < n; i++) forward((Transformer*)hs[i], token[i], pos[i]);
}

int refq8_argmax(float* p, int n) { return sample_argmax(p, n); }
int refq8_max_threads(void) { return omp_get_max_threads(); }
