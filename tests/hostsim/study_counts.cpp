// tests/hostsim/study_counts.cpp -- TEST/STUDY ONLY: per-query work counters of the product
// traversal (box tests, point evaluations), to study load imbalance on the CPU build box.
#define PCNN_HOST_COUNTERS 1
#include "hostsim.cpp"
namespace pcnn { thread_local long g_cnt_box = 0, g_cnt_pt = 0; }
extern "C" int hostsim_knn_counts(void* hp, const float* q, int64_t nq, int stride, int leaf_log,
                                  int64_t* box, int64_t* pt) {
    const HostIndex* h = (const HostIndex*)hp;
    for (int64_t i = 0; i < nq; i++) {
        pcnn::g_cnt_box = 0; pcnn::g_cnt_pt = 0;
        TopK<8> tk;
        knn_thread<8>(h->v, q[i * stride], q[i * stride + 1], q[i * stride + 2], leaf_log, tk);
        box[i] = pcnn::g_cnt_box; pt[i] = pcnn::g_cnt_pt;
    }
    return 0;
}
