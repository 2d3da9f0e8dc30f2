/* Plain C caller of the placement library: the C ABI needs no Python, no torch and no CUDA
 * headers on the caller's side.  It places a handful of pods on a small node table through the
 * host-pointer session (kp_session_snapshot, kp_session_decide per pod, then kp_session_greedy) and prints the results, one
 * line per pod; tests/test_c_abi_example_gpu.py compiles it with gcc, runs it on the GPU box and
 * compares the lines with the oracle.
 *
 *   gcc -O2 -I include examples/c_abi_example.c -L k8s_llm_scheduler_b200 -lk8s_place_b200 \
 *       -Wl,-rpath,$PWD/k8s_llm_scheduler_b200 -o /tmp/c_abi_example && /tmp/c_abi_example
 *
 * Node rows are the reference's NodeMetrics in wire units (scheduler.py:72-84): milli-cores,
 * bytes, pod counts, usage per cent as (pod_count / max_pods) * 50 (scheduler.py:150-151).
 */
#include <stdint.h>
#include <stdio.h>

#include "k8s_place_b200.h"

#define N_NODES 6
#define N_PODS 8

static int fail(const char *what, int status)
{
    fprintf(stderr, "%s: status %d (%s): %s\n", what, status, kp_status_string(status), kp_last_error());
    return 1;
}

int main(void)
{
    int64_t cpu[N_NODES] = {4000, 16000, 32000, 32000, 96000, 8000};
    int64_t mem[N_NODES] = {16ll << 30, 64ll << 30, 128ll << 30, 128ll << 30, 768ll << 30, 15ll << 30};
    int64_t pc[N_NODES] = {3, 20, 55, 12, 100, 8};
    int64_t mp[N_NODES] = {30, 58, 110, 110, 250, 110};
    uint8_t ready[N_NODES] = {1, 1, 1, 0, 1, 1};
    double cu[N_NODES], mu[N_NODES];
    for (int i = 0; i < N_NODES; ++i) cu[i] = mu[i] = ((double)pc[i] / (double)mp[i]) * 50;
    int64_t req_cpu[N_PODS] = {250, 500, 100, 20000, 64000, 3000, 1, 99000};
    int64_t req_mem[N_PODS] = {256ll << 20, 512ll << 20, 128ll << 20, 40ll << 30, 500ll << 30, 10ll << 30, 1, 1ll << 30};

    kp_node_table nodes = {N_NODES, cpu, mem, pc, mp, ready, cu, mu};
    kp_pod_table pods = {N_PODS, req_cpu, req_mem, NULL};
    kp_session *s = NULL;
    int st = kp_session_create(0, 64, 64, &s);
    if (st != KP_OK) return fail("kp_session_create", st);

    int64_t idx[N_PODS], final_pc[N_NODES];
    double score[N_PODS], final_cu[N_NODES], final_mu[N_NODES];
    for (int strategy = 0; strategy < 3; ++strategy) {
        st = kp_session_snapshot(s, &nodes, &pods, strategy, KP_FILTER_CAPACITY, idx, score);
        if (st != KP_OK) return fail("kp_session_snapshot", st);
        for (int j = 0; j < N_PODS; ++j)
            printf("snapshot strategy=%d pod=%d node=%lld score=%a\n", strategy, j, (long long)idx[j], score[j]);
        /* the reference's own call shape: one decision per pod over the node list (scheduler.py:598 -> :521-559) */
        for (int j = 0; j < N_PODS; ++j) {
            int64_t one_idx;
            double one_score;
            st = kp_session_decide(s, &nodes, strategy, KP_FILTER_CAPACITY, req_cpu[j], req_mem[j], &one_idx, &one_score);
            if (st != KP_OK) return fail("kp_session_decide", st);
            printf("decide strategy=%d pod=%d node=%lld score=%a\n", strategy, j, (long long)one_idx, one_score);
        }
        st = kp_session_greedy(s, NULL, &pods, strategy, KP_FILTER_CAPACITY, idx, score, final_pc, final_cu, final_mu);
        if (st != KP_OK) return fail("kp_session_greedy", st);
        for (int j = 0; j < N_PODS; ++j)
            printf("greedy strategy=%d pod=%d node=%lld score=%a\n", strategy, j, (long long)idx[j], score[j]);
        for (int i = 0; i < N_NODES; ++i)
            printf("greedy strategy=%d final node=%d pod_count=%lld cpu_usage=%a\n", strategy, i, (long long)final_pc[i], final_cu[i]);
    }
    /* errors come back as status codes with a message, never as exceptions */
    st = kp_session_snapshot(s, &nodes, &pods, 0, 7, idx, score);
    printf("bad filter -> status=%d (%s)\n", st, kp_status_string(st));
    kp_session_destroy(s);
    return 0;
}
