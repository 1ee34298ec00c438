/* Plain C99 client of include/spif_b200.h: what a C / cgo / JNI host sees.  Host-only entry points are exercised for real
 * (model file, model split, placement, balancer state, sizes, error strings); compute entry points are only checked to fail
 * loudly without a device.  Built and run by tests/test_c_client.py:  gcc -std=c99 host_abi.c -lspif_b200  */
#include "spif_b200.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define CHECK(cond)                                                         \
    do {                                                                    \
        if (!(cond)) {                                                      \
            fprintf(stderr, "%s:%d: CHECK(%s) failed\n", __FILE__, __LINE__, #cond); \
            return 1;                                                       \
        }                                                                   \
    } while (0)

int main(int argc, char ** argv) {
    CHECK(argc == 3); /* argv[1] = model GGUF (tests/golden/model_tiny_ref.gguf), argv[2] = scratch path for a model split */
    CHECK(spif_abi_version() == SPIF_B200_ABI_VERSION);
    CHECK(spif_row_size(SPIF_TYPE_F16, 5120) == 10240 && spif_row_size(SPIF_TYPE_Q8_0, 5120) == 5440 &&
          spif_row_size(SPIF_TYPE_Q4_0, 5120) == 2880);
    CHECK(spif_workspace_size(5120, 13824, 1) > 0 && spif_workspace_size(5120, 13824, 1) <= spif_noalloc_max_bytes());
    CHECK(strstr(spif_error_string(SPIF_E_WORKSPACE), "workspace") != NULL);

    /* model file */
    spif_gguf * g = NULL;
    CHECK(spif_gguf_open(argv[1], &g) == 0 && g != NULL);
    char    arch[64];
    int32_t n_layer = 0, n_embd = 0, n_ff = 0;
    float   eps = 0.f;
    CHECK(spif_gguf_hparams(g, arch, (int) sizeof(arch), &n_layer, &n_embd, &n_ff, &eps) == 0);
    CHECK(strcmp(arch, "prosparse-llama") == 0 && n_layer == 2 && n_embd == 64 && n_ff == 128 && eps > 0.f);
    int32_t rank[2];
    CHECK(spif_gguf_pred_rank(g, rank, n_layer) == 0 && rank[0] == 32 && rank[1] == 32);
    spif_gguf_tensor t;
    CHECK(spif_gguf_find(g, "blk.1.ffn_down.weight", &t) == 0);
    CHECK(t.type == SPIF_TYPE_F16 && t.n_dims == 2 && t.ne[0] == n_embd && t.ne[1] == n_ff); /* one row per neuron */
    unsigned char * all  = (unsigned char *) malloc((size_t) t.nbytes);
    unsigned char * part = (unsigned char *) malloc((size_t) t.nbytes);
    CHECK(all && part);
    CHECK(spif_gguf_read(g, &t, 0, n_ff, all, (size_t) t.nbytes) == 0);
    const size_t rb = spif_row_size(t.type, t.ne[0]);
    CHECK(spif_gguf_read(g, &t, 16, 32, part, 32 * rb) == 0 && memcmp(part, all + 16 * rb, 32 * rb) == 0);
    int32_t idx[3] = { 127, 0, 64 };
    CHECK(spif_gguf_gather(g, &t, idx, 3, part, 3 * rb) == 0 && memcmp(part, all + 127 * rb, rb) == 0 &&
          memcmp(part + rb, all, rb) == 0 && memcmp(part + 2 * rb, all + 64 * rb, rb) == 0);
    CHECK(spif_gguf_read(g, &t, 120, 16, part, 16 * rb) == SPIF_E_SHAPE);    /* past the end */
    CHECK(spif_gguf_read(g, &t, 0, 8, part, 8 * rb - 1) == SPIF_E_WORKSPACE); /* destination too small */
    CHECK(spif_gguf_find(g, "blk.7.ffn_up.weight", &t) == SPIF_E_SHAPE);
    CHECK(spif_gguf_n_tensors(g) == 16);
    spif_gguf_close(g);

    /* model split: write -> open -> the load-time pipeline (budget per layer, row reorder) */
    enum { L = 2, F = 128, G = 16 };
    float   pattern[L] = { 0.25f, 0.75f };
    int32_t perms[L][F];
    for (int l = 0; l < L; ++l) {
        for (int i = 0; i < F; ++i) {
            perms[l][i] = (i * 37 + 11 * l) % F; /* 37 is coprime to 128: a permutation */
        }
    }
    CHECK(spif_model_split_write(argv[2], G, pattern, L, &perms[0][0], F) == 0);
    spif_model_split * ms = NULL;
    CHECK(spif_model_split_open(argv[2], &ms) == 0);
    int32_t gs = 0, nl = 0, nf = 0;
    CHECK(spif_model_split_info(ms, &gs, &nl, &nf) == 0 && gs == G && nl == L && nf == F);
    float   pat2[L];
    int32_t perm2[F];
    CHECK(spif_model_split_pattern(ms, pat2, L) == 0 && pat2[0] == 0.25f && pat2[1] == 0.75f);
    CHECK(spif_model_split_perm(ms, 1, perm2, F) == 0 && memcmp(perm2, perms[1], sizeof(perm2)) == 0);
    spif_model_split_close(ms);
    int32_t groups[L];
    CHECK(spif_cache_group_budget((int64_t) 3 * 4 * G * rb, (int64_t) G * rb, 1, pattern, L, F / G, groups) == 0);
    CHECK(groups[0] >= 0 && groups[1] >= groups[0] && groups[1] <= F / G);
    CHECK(spif_reorder_rows(all, part, perm2, F, (int64_t) rb) == 0 && memcmp(part, all + (size_t) perm2[0] * rb, rb) == 0);

    /* placement + balancer state machine, host only */
    float   heat[8] = { 8, 7, 6, 5, 4, 3, 2, 1 };
    int32_t shard[8];
    CHECK(spif_initial_placement(heat, 8, 2, shard) == 0);
    int n0 = 0;
    for (int i = 0; i < 8; ++i) {
        n0 += shard[i] == 0;
    }
    CHECK(n0 == 4);
    spif_balancer * b = NULL;
    CHECK(spif_balancer_create(1, 8, 2, 0, shard, 0.9f, 0.05f, 0.02f, &b) == 0 && b != NULL);
    int32_t slots[8]; /* [n_shards][n_slot] -> group id */
    CHECK(spif_balancer_slots(b, 0, slots) == 0);
    for (int s = 0; s < 2; ++s) {
        for (int k = 0; k < 4; ++k) {
            CHECK(slots[s * 4 + k] >= 0 && slots[s * 4 + k] < 8 && shard[slots[s * 4 + k]] == s);
        }
    }
    float   lam = 0.f;
    int64_t swaps = -1;
    int     pending = -1;
    CHECK(spif_balancer_state(b, &lam, &swaps, &pending) == 0 && lam == 0.9f && swaps == 0 && pending == 0);
    spif_balancer_destroy(b);

    /* compute entry points: an error code without a device, never a host computation */
    int sms = 0;
    if (spif_device_sm_count(&sms) != 0) {
        float buf[64] = { 0 };
        CHECK(spif_fatrelu(buf, buf, 64, 0.01f, NULL) != 0);
    }
    free(all);
    free(part);
    printf("host_abi ok\n");
    return 0;
}
