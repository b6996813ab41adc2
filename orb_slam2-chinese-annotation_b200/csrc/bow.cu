// Bag-of-words transform on the device: the vocabulary-tree descent behind Frame::ComputeBoW /
// KeyFrame::ComputeBoW (reference src/Frame.cc:439-449, src/KeyFrame.cc:60-71), i.e.
// DBoW2::TemplatedVocabulary<FORB>::transform(feature, word_id, weight, nid, levelsup)
// (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1239-1283) for every descriptor of a frame:
// from the root, move to the child whose descriptor has the smallest Hamming distance (strict <,
// the first child wins ties) until a leaf is reached; report the leaf's word id and the node passed
// at level L - levelsup.  Pure XOR/POPC work on 32-byte rows; the tree (35 MB for the 10^6-word ORB
// vocabulary) lives in L2 after the first frames.
//
// Device layout: nodes are renumbered so that the children of a node are CONSECUTIVE slots (slot s
// is internal node s + 1; the root is internal node 0): one group of lanes reads a node's k child
// descriptors as one contiguous k x 32 B block.  d_first/d_cnt give the child range of an internal
// node, d_orig maps back to the caller's node ids, d_word holds the word id of leaves (-1 otherwise).
#include <algorithm>
#include <cstring>
#include <mutex>
#include "orb_internal.h"

struct orb_vocabulary {
    int device = 0;
    int n_nodes = 0, n_words = 0, levels = 0, max_k = 0, max_features = 0;
    cudaStream_t stream = nullptr;
    uint4* d_desc = nullptr;      // [n_nodes - 1][2]  child descriptors by slot
    int2* d_range = nullptr;      // [n_nodes] (first slot, count) of an internal node's children
    int* d_orig = nullptr;        // [n_nodes] internal -> caller node id
    int* d_word = nullptr;        // [n_nodes] internal -> word id or -1
    uint8_t* d_q = nullptr;       // [max_features][32] staging for host descriptors
    int* d_out = nullptr;         // [2][max_features] word id, node id
    long long launches = 0;
    // orb_bow_transform stages through d_q / d_out on the one stream: calls on one handle are serialised here, because the
    // reference transforms on one shared vocabulary from two threads (Frame::ComputeBoW on Tracking, KeyFrame::ComputeBoW on
    // LocalMapping).  The tree itself is read-only; orb_bow_transform_device leaves ordering to the caller's streams.
    std::mutex host_call;
};

namespace {

// One group of G lanes (G = 16 or 32, G >= largest child count) per descriptor.
template <int G>
__global__ void __launch_bounds__(256) k_bow_transform(const uint4* __restrict__ q, int n, const uint4* __restrict__ desc, const int2* __restrict__ range,
                                                        const int* __restrict__ orig, const int* __restrict__ word, int nid_level,
                                                        int* __restrict__ out_word, int* __restrict__ out_node) {
    const int lane = threadIdx.x & (G - 1);
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) / G;
    if (i >= n) return;   // whole groups leave together: n is tested per group
    const unsigned gmask = G == 32 ? 0xffffffffu : (0xffffu << (threadIdx.x & 16));
    const uint4 q0 = __ldg(q + 2 * (size_t)i), q1 = __ldg(q + 2 * (size_t)i + 1);
    int node = 0, level = 0, nid = nid_level <= 0 ? 0 : -1;
    int2 r = __ldg(range + 0);
    while (r.y > 0) {
        ++level;
        unsigned key = 0xffffffffu;
        if (lane < r.y) {
            const uint4* c = desc + 2 * (size_t)(r.x + lane);
            const int d = orb_hamming256(q0, q1, __ldg(c), __ldg(c + 1));
            key = ((unsigned)d << 8) | (unsigned)lane;        // smallest distance, then first child
        }
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) key = min(key, __shfl_xor_sync(gmask, key, o, G));
        node = r.x + (int)(key & 0xffu) + 1;
        if (level == nid_level) nid = node;
        r = __ldg(range + node);
    }
    if (lane == 0) {
        out_word[i] = __ldg(word + node);
        out_node[i] = __ldg(orig + (nid < 0 ? node : nid));   // a leaf above level L - levelsup stands for itself
    }
}

}  // namespace

extern "C" {

int orb_vocabulary_create(int levels, int n_nodes, const int32_t* parent, const uint8_t* is_leaf, const uint8_t* desc, int max_features,
                          int device, orb_vocabulary** out) {
    if (!out) return ORB_ERR_BAD_ARG;
    *out = nullptr;
    if (n_nodes < 2 || levels < 1 || !parent || !is_leaf || !desc || max_features < 1) { orb_set_error("bad vocabulary arguments"); return ORB_ERR_BAD_ARG; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        orb_set_error("no CUDA device: this library has no CPU fallback");
        return ORB_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) { orb_set_error("bad device index"); return ORB_ERR_BAD_ARG; }
    // children lists in file order (loadFromTextFile appends node ids as they are read: TemplatedVocabulary.h:1409-1416)
    for (int i = 1; i < n_nodes; ++i)
        if (parent[i] < 0 || parent[i] >= i) { orb_set_error("vocabulary: a node must come after its parent"); return ORB_ERR_BAD_ARG; }
    // internal numbering: breadth-first from the root, children of one node consecutive
    std::vector<int> order(1, 0), inner(n_nodes, -1);
    std::vector<std::vector<int>> kids(n_nodes);
    for (int i = 1; i < n_nodes; ++i) kids[parent[i]].push_back(i);
    order.reserve(n_nodes); inner[0] = 0;
    std::vector<int2> range(n_nodes);
    int max_k = 0;
    for (size_t h = 0; h < order.size(); ++h) {
        const int p = order[h];
        range[h] = make_int2((int)order.size() - 1, (int)kids[p].size());
        max_k = std::max(max_k, (int)kids[p].size());
        for (int c : kids[p]) { inner[c] = (int)order.size(); order.push_back(c); }
    }
    if ((int)order.size() != n_nodes) { orb_set_error("vocabulary: unreachable nodes"); return ORB_ERR_BAD_ARG; }
    if (max_k > 32) { orb_set_error("vocabulary: more than 32 children per node"); return ORB_ERR_BAD_ARG; }
    std::vector<int> word(n_nodes, -1), wid_of(n_nodes, -1);
    int n_words = 0;
    for (int i = 1; i < n_nodes; ++i) if (is_leaf[i]) wid_of[i] = n_words++;      // word ids in file order (:1432-1440)
    std::vector<uint8_t> slot_desc((size_t)(n_nodes - 1) * 32);
    for (int h = 0; h < n_nodes; ++h) {
        word[h] = wid_of[order[h]];
        if (is_leaf[order[h]] && range[h].y != 0) { orb_set_error("vocabulary: a leaf has children"); return ORB_ERR_BAD_ARG; }
        if (!is_leaf[order[h]] && range[h].y == 0 && h != 0) { orb_set_error("vocabulary: an inner node has no children"); return ORB_ERR_BAD_ARG; }
        if (h > 0) memcpy(&slot_desc[(size_t)(h - 1) * 32], desc + (size_t)order[h] * 32, 32);
    }
    ORB_CUDA(cudaSetDevice(device));
    orb_vocabulary* v = new orb_vocabulary();
    v->device = device; v->n_nodes = n_nodes; v->n_words = n_words; v->levels = levels; v->max_k = max_k; v->max_features = max_features;
    if (cudaStreamCreateWithFlags(&v->stream, cudaStreamNonBlocking) != cudaSuccess ||
        orb_dev_alloc((void**)&v->d_desc, slot_desc.size()) != cudaSuccess || orb_dev_alloc((void**)&v->d_range, sizeof(int2) * n_nodes) != cudaSuccess ||
        orb_dev_alloc((void**)&v->d_orig, 4 * (size_t)n_nodes) != cudaSuccess || orb_dev_alloc((void**)&v->d_word, 4 * (size_t)n_nodes) != cudaSuccess ||
        orb_dev_alloc((void**)&v->d_q, (size_t)max_features * 32) != cudaSuccess || orb_dev_alloc((void**)&v->d_out, (size_t)max_features * 8) != cudaSuccess ||
        cudaMemcpy(v->d_desc, slot_desc.data(), slot_desc.size(), cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMemcpy(v->d_range, range.data(), sizeof(int2) * n_nodes, cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMemcpy(v->d_orig, order.data(), 4 * (size_t)n_nodes, cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMemcpy(v->d_word, word.data(), 4 * (size_t)n_nodes, cudaMemcpyHostToDevice) != cudaSuccess) {
        orb_set_error(std::string("vocabulary allocation: ") + cudaGetErrorString(cudaGetLastError()));
        orb_vocabulary_destroy(v);
        return ORB_ERR_CUDA;
    }
    *out = v;
    return ORB_OK;
}

void orb_vocabulary_destroy(orb_vocabulary* v) {
    if (!v) return;
    cudaSetDevice(v->device);
    if (v->stream) cudaStreamSynchronize(v->stream);
    orb_dev_free(v->d_desc); orb_dev_free(v->d_range); orb_dev_free(v->d_orig); orb_dev_free(v->d_word); orb_dev_free(v->d_q); orb_dev_free(v->d_out);
    if (v->stream) cudaStreamDestroy(v->stream);
    delete v;
}

int orb_vocabulary_info(const orb_vocabulary* v, int32_t* n_nodes, int32_t* n_words, int32_t* max_children) {
    if (!v) return ORB_ERR_BAD_ARG;
    if (n_nodes) *n_nodes = v->n_nodes;
    if (n_words) *n_words = v->n_words;
    if (max_children) *max_children = v->max_k;
    return ORB_OK;
}
void* orb_vocabulary_stream(const orb_vocabulary* v) { return v ? (void*)v->stream : nullptr; }
long long orb_vocabulary_launch_count(const orb_vocabulary* v) { return v ? v->launches : 0; }

int orb_bow_transform_device(orb_vocabulary* v, const uint8_t* d_desc, int n, int levelsup, int32_t* d_word_id, int32_t* d_node_id) {
    ORB_NVTX("orb:bow_transform_device");
    if (!v || n < 0 || (n > 0 && (!d_desc || !d_word_id || !d_node_id))) return ORB_ERR_BAD_ARG;
    if (((uintptr_t)d_desc & 15) != 0) { orb_set_error("descriptors must be 16-byte aligned"); return ORB_ERR_BAD_ARG; }
    if (n == 0) return ORB_OK;
    ORB_CUDA(cudaSetDevice(v->device));
    const int nid_level = v->levels - levelsup;
    if (v->max_k <= 16) {
        const int per_cta = 256 / 16;
        k_bow_transform<16><<<(n + per_cta - 1) / per_cta, 256, 0, v->stream>>>((const uint4*)d_desc, n, v->d_desc, v->d_range, v->d_orig, v->d_word, nid_level, d_word_id, d_node_id);
    } else {
        const int per_cta = 256 / 32;
        k_bow_transform<32><<<(n + per_cta - 1) / per_cta, 256, 0, v->stream>>>((const uint4*)d_desc, n, v->d_desc, v->d_range, v->d_orig, v->d_word, nid_level, d_word_id, d_node_id);
    }
    v->launches++;
    ORB_CUDA(cudaGetLastError());
    return ORB_OK;
}

int orb_bow_transform(orb_vocabulary* v, const uint8_t* desc, int n, int levelsup, int32_t* word_id, int32_t* node_id) {
    ORB_NVTX("orb:bow_transform");
    if (!v || n < 0 || (n > 0 && (!desc || !word_id || !node_id))) return ORB_ERR_BAD_ARG;
    if (n > v->max_features) { orb_set_error("more descriptors than the vocabulary handle was created for"); return ORB_ERR_CAPACITY; }
    if (n == 0) return ORB_OK;
    std::lock_guard<std::mutex> serialise(v->host_call);
    ORB_CUDA(cudaSetDevice(v->device));
    ORB_CUDA(cudaMemcpyAsync(v->d_q, desc, (size_t)n * 32, cudaMemcpyHostToDevice, v->stream));
    const int st = orb_bow_transform_device(v, v->d_q, n, levelsup, v->d_out, v->d_out + v->max_features);
    if (st != ORB_OK) return st;
    ORB_CUDA(cudaMemcpyAsync(word_id, v->d_out, (size_t)n * 4, cudaMemcpyDeviceToHost, v->stream));
    ORB_CUDA(cudaMemcpyAsync(node_id, v->d_out + v->max_features, (size_t)n * 4, cudaMemcpyDeviceToHost, v->stream));
    ORB_CUDA(cudaStreamSynchronize(v->stream));
    return ORB_OK;
}

int orb_vocabulary_sync(orb_vocabulary* v) {
    if (!v) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(v->device));
    ORB_CUDA(cudaStreamSynchronize(v->stream));
    return ORB_OK;
}

}  // extern "C"
