// bow.cu — the DBoW2 vocabulary-tree transform behind Frame::ComputeBoW / KeyFrame::ComputeBoW
// (reference src/Frame.cc:905-922: mpORBvocabulary->transform(vCurrentDesc, mBowVec, mFeatVec, 4)) on the device
// (SURVEY.md 8f rank 3).
//
// Vocabulary (orbx_vocabulary): the tree of Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h resident in HBM as CSR —
// children of node i = child_idx[child_start[i] .. child_start[i+1]) in push_back order (loadFromTextFile
// :1338-1415), node descriptors 32 B each, leaf -> (word id, weight).  ORBvoc (k = 10, L = 6) is 1.1 M nodes = 36 MB
// of descriptors; the top levels stay in L2.
// bow_transform_kernel: one warp per feature descends the tree (TemplatedVocabulary::transform :1218-1258): at
// every level the lanes take the children of the current node, FORB::distance (FORB.cpp:81-100) = XOR + POPC, warp
// arg-min over (distance, child position) = "first least child wins"; the node passed at level L - levelsup is the
// FeatureVector key.  Latency-bound pointer chase of L dependent steps; thousands of features hide it.
// Host side: BowVector / FeatureVector assembly in std::map order (transform :1127-1196, BowVector.cpp:34-84) in
// double — a few thousand scalar steps.
#include "match_internal.h"
#include <climits>
#include <cmath>
#include <cstring>
#include <map>
#include <vector>

struct orbx_vocabulary {
    int device; int k, L; int nnodes;
    int* d_child_start; int* d_child_idx; uint4* d_desc; double* d_weight; int* d_word;
};

#define BOW_WARPS 8

__global__ void __launch_bounds__(BOW_WARPS * 32)
bow_transform_kernel(const int* __restrict__ child_start, const int* __restrict__ child_idx, const uint4* __restrict__ node_desc,
                     const double* __restrict__ node_weight, const int* __restrict__ node_word, int nid_level,
                     int n, const uint4* __restrict__ desc, int* __restrict__ word, double* __restrict__ weight, int* __restrict__ node)
{
    const int f = blockIdx.x * BOW_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (f >= n) return;
    const uint4 a0 = desc[2 * (size_t)f], a1 = desc[2 * (size_t)f + 1];
    int nid = 0, final_id = 0, level = 0;
    int b = child_start[0], e = child_start[1];
    do {
        ++level;
        int bestD = INT_MAX, bestC = INT_MAX;
        for (int c = b + lane; c < e; c += 32) {
            const int id = child_idx[c];
            const int d = ham256(a0, a1, node_desc[2 * (size_t)id], node_desc[2 * (size_t)id + 1]);
            if (d < bestD) { bestD = d; bestC = c; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const int od = __shfl_xor_sync(0xffffffffu, bestD, o), oc = __shfl_xor_sync(0xffffffffu, bestC, o);
            if (od < bestD || (od == bestD && oc < bestC)) { bestD = od; bestC = oc; }
        }
        final_id = child_idx[bestC];
        if (level == nid_level) nid = final_id;
        b = child_start[final_id]; e = child_start[final_id + 1];
    } while (b != e);                                                          // !isLeaf()
    if (lane == 0) { word[f] = node_word[final_id]; weight[f] = node_weight[final_id]; node[f] = nid; }
}

extern "C" int orbx_vocabulary_create(int device, int k, int L, int nrows, const int32_t* parent, const uint8_t* is_leaf,
                                      const uint8_t* desc, const double* weight, orbx_vocabulary** out)
{
    if (!out) { orbx_set_error("out is NULL"); return ORBX_ERR_ARG; }
    *out = nullptr;
    if (nrows < 1 || !parent || !is_leaf || !desc || !weight || L < 1 || L > 10 || k < 0 || k > 20) {   // :1359
        orbx_set_error("bad vocabulary arguments"); return ORBX_ERR_ARG;
    }
    if (device < 0 || device >= orbx_device_count()) { orbx_set_error("CUDA device %d not available; no CPU path", device); return ORBX_ERR_CUDA; }
    const int nn = nrows + 1;
    // children in push_back order = file order (counting sort by parent keeps it), word ids number the leaves in file order
    std::vector<int> cs((size_t)nn + 1, 0), ci((size_t)nrows), word((size_t)nn, 0);
    std::vector<double> w((size_t)nn, 0.0);
    for (int r = 0; r < nrows; r++) {
        if (parent[r] < 0 || parent[r] > r) { orbx_set_error("vocabulary row %d: parent %d is not an earlier node", r, parent[r]); return ORBX_ERR_ARG; }
        cs[(size_t)parent[r] + 1]++;
    }
    for (int i = 0; i < nn; i++) cs[(size_t)i + 1] += cs[(size_t)i];
    {
        std::vector<int> fill(cs.begin(), cs.end() - 1);
        int wid = 0;
        for (int r = 0; r < nrows; r++) {
            ci[(size_t)fill[(size_t)parent[r]]++] = r + 1;
            w[(size_t)r + 1] = weight[r];
            if (is_leaf[r]) word[(size_t)r + 1] = wid++;
        }
    }
    for (int r = 0; r < nrows; r++) {
        const bool has_children = cs[(size_t)r + 2] != cs[(size_t)r + 1];
        if (has_children == (is_leaf[r] != 0)) { orbx_set_error("vocabulary row %d: leaf flag contradicts the tree", r); return ORBX_ERR_ARG; }
    }
    if (cs[1] == cs[0]) { orbx_set_error("vocabulary root has no children"); return ORBX_ERR_ARG; }
    orbx_vocabulary* v = new orbx_vocabulary();
    memset(v, 0, sizeof(*v));
    v->device = device; v->k = k; v->L = L; v->nnodes = nn;
    cudaError_t ce = cudaSetDevice(device);
    if (ce == cudaSuccess) ce = cudaMalloc((void**)&v->d_child_start, sizeof(int) * ((size_t)nn + 1));
    if (ce == cudaSuccess) ce = cudaMalloc((void**)&v->d_child_idx, sizeof(int) * (size_t)nrows);
    if (ce == cudaSuccess) ce = cudaMalloc((void**)&v->d_desc, 32 * (size_t)nn);
    if (ce == cudaSuccess) ce = cudaMalloc((void**)&v->d_weight, sizeof(double) * (size_t)nn);
    if (ce == cudaSuccess) ce = cudaMalloc((void**)&v->d_word, sizeof(int) * (size_t)nn);
    if (ce == cudaSuccess) ce = cudaMemcpy(v->d_child_start, cs.data(), sizeof(int) * ((size_t)nn + 1), cudaMemcpyHostToDevice);
    if (ce == cudaSuccess) ce = cudaMemcpy(v->d_child_idx, ci.data(), sizeof(int) * (size_t)nrows, cudaMemcpyHostToDevice);
    if (ce == cudaSuccess) ce = cudaMemset(v->d_desc, 0, 32);
    if (ce == cudaSuccess) ce = cudaMemcpy((char*)v->d_desc + 32, desc, 32 * (size_t)nrows, cudaMemcpyHostToDevice);
    if (ce == cudaSuccess) ce = cudaMemcpy(v->d_weight, w.data(), sizeof(double) * (size_t)nn, cudaMemcpyHostToDevice);
    if (ce == cudaSuccess) ce = cudaMemcpy(v->d_word, word.data(), sizeof(int) * (size_t)nn, cudaMemcpyHostToDevice);
    if (ce != cudaSuccess) {
        orbx_set_error("vocabulary upload failed: %s", cudaGetErrorString(ce));
        cudaFree(v->d_child_start); cudaFree(v->d_child_idx); cudaFree(v->d_desc); cudaFree(v->d_weight); cudaFree(v->d_word);
        delete v;
        return ORBX_ERR_CUDA;
    }
    *out = v;
    return ORBX_OK;
}

extern "C" void orbx_vocabulary_destroy(orbx_vocabulary* v)
{
    if (!v) return;
    cudaSetDevice(v->device);
    cudaFree(v->d_child_start); cudaFree(v->d_child_idx); cudaFree(v->d_desc); cudaFree(v->d_weight); cudaFree(v->d_word);
    delete v;
}

extern "C" int orbx_bow_transform(const orbx_vocabulary* v, const uint8_t* desc, int n, int levelsup,
                                  int32_t* word, double* weight, int32_t* node)
{
    if (!v || n < 0 || (n > 0 && (!desc || !word || !weight || !node))) { orbx_set_error("bad bow_transform arguments"); return ORBX_ERR_ARG; }
    if (n == 0) return ORBX_OK;
    OrbxWorkspace& W = orbx_workspace();
    int rc = W.bind(v->device);
    if (rc != ORBX_OK) return rc;
    OrbxCarver c;
    const size_t o_desc = c.take((size_t)n * 32);
    const size_t in_bytes = c.off;
    const size_t o_weight = c.take((size_t)n * 8), o_word = c.take((size_t)n * 4), o_node = c.take((size_t)n * 4);
    rc = W.reserve(c.off, c.off);
    if (rc != ORBX_OK) return rc;
    memcpy(W.h + o_desc, desc, (size_t)n * 32);
    CUDA_TRYW(cudaMemcpyAsync(W.d, W.h, in_bytes, cudaMemcpyHostToDevice, W.s));
    bow_transform_kernel<<<(n + BOW_WARPS - 1) / BOW_WARPS, BOW_WARPS * 32, 0, W.s>>>(
        v->d_child_start, v->d_child_idx, v->d_desc, v->d_weight, v->d_word, v->L - levelsup, n,
        (const uint4*)(W.d + o_desc), (int*)(W.d + o_word), (double*)(W.d + o_weight), (int*)(W.d + o_node));
    CUDA_TRYW(cudaGetLastError());
    CUDA_TRYW(cudaMemcpyAsync(W.h + o_weight, W.d + o_weight, c.off - o_weight, cudaMemcpyDeviceToHost, W.s));
    CUDA_TRYW(cudaStreamSynchronize(W.s));
    memcpy(weight, W.h + o_weight, (size_t)n * 8);
    memcpy(word, W.h + o_word, (size_t)n * 4);
    memcpy(node, W.h + o_node, (size_t)n * 4);
    return ORBX_OK;
}

extern "C" int orbx_compute_bow(const orbx_vocabulary* v, const uint8_t* desc, int n, int levelsup,
                                int32_t* bow_id, double* bow_val, int* nbow,
                                int32_t* fv_node, int32_t* fv_start, int32_t* fv_idx, int* nfv)
{
    if (!v || n < 0 || !nbow || !nfv || !fv_start || (n > 0 && (!desc || !bow_id || !bow_val || !fv_node || !fv_idx))) {
        orbx_set_error("bad compute_bow arguments"); return ORBX_ERR_ARG;
    }
    *nbow = 0; *nfv = 0; fv_start[0] = 0;
    if (n == 0) return ORBX_OK;
    std::vector<int32_t> word((size_t)n), node((size_t)n);
    std::vector<double> weight((size_t)n);
    const int rc = orbx_bow_transform(v, desc, n, levelsup, word.data(), weight.data(), node.data());
    if (rc != ORBX_OK) return rc;
    std::map<int32_t, double> bv;                          // BowVector: std::map<WordId, WordValue>
    std::map<int32_t, std::vector<int32_t>> fv;            // FeatureVector: std::map<NodeId, std::vector<unsigned int>>
    for (int i = 0; i < n; i++) {
        if (!(weight[(size_t)i] > 0)) continue;            // stopped word, TemplatedVocabulary.h:1157
        bv[word[(size_t)i]] += weight[(size_t)i];          // addWeight, BowVector.cpp:34-46
        fv[node[(size_t)i]].push_back(i);                  // addFeature
    }
    double norm = 0.0;                                     // normalize(L1), BowVector.cpp:62-84
    for (auto& kv : bv) norm += fabs(kv.second);
    int m = 0;
    for (auto& kv : bv) { bow_id[m] = kv.first; bow_val[m] = norm > 0.0 ? kv.second / norm : kv.second; m++; }
    *nbow = m;
    int k = 0, pos = 0;
    for (auto& kv : fv) {
        fv_node[k] = kv.first; fv_start[k] = pos;
        for (int32_t i : kv.second) fv_idx[pos++] = i;
        k++;
    }
    fv_start[k] = pos;
    *nfv = k;
    return ORBX_OK;
}
