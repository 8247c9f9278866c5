// Batch prediction with decision-path models:
//   DecisionTreeModel.predictClassProb  (tree/DecisionTreeModel.java:56-93): the first decision path, in
//     list order, whose predicates all hold for the row wins;
//   EnsemblePredictiveModel.predict     (model/EnsemblePredictiveModel.java:69-110): weighted vote of the
//     member models, highest total first, ties in the iteration order of the vote map.
// The host turns every model into admission masks: for (attribute a, bin b) a bit set over the model's
// paths, bit p set when path p admits bin b of attribute a (all of the path's predicates on a hold).
// A row matches the paths in  AND_a mask[a][code_a(row)];  the lowest set bit is the first matching path.
// Included by avt_kernels.cuh.
#pragma once

namespace avt {

struct PredictArgs {
    const void* const* cols;     // [A] bin-code columns of the loaded data set
    const int32_t* wide;         // [A]
    const int32_t* bin_off;      // [A+1]
    const uint32_t* masks;       // model m: [TB][words_m] at mask_off[m]
    const int64_t* mask_off;     // [M]
    const int32_t* n_paths;      // [M]
    const int32_t* path_off;     // [M] offset of model m's paths in path_class
    const uint8_t* path_class;   // output class code of every path
    const double* weights;       // [M]
    const uint8_t* class_rank;   // [n_classes] position in the vote map's iteration order
    uint8_t* out;                // [N] predicted class code, 0xFF = no decision path matched in some model
    long long N;
    int A, M, n_classes;
};

__global__ void __launch_bounds__(256) predict_kernel(const PredictArgs a) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < a.N; r += stride) {
        double votes[kMaxC];
        for (int c = 0; c < a.n_classes; ++c) votes[c] = 0.0;
        bool voted[kMaxC];
        for (int c = 0; c < a.n_classes; ++c) voted[c] = false;
        bool unmatched = false;
        int last = 0;
        for (int m = 0; m < a.M; ++m) {
            const int P = a.n_paths[m], W = (P + 31) >> 5;
            const uint32_t* mk = a.masks + a.mask_off[m];
            int path = -1;
            for (int w = 0; w < W && path < 0; ++w) {
                uint32_t acc = 0xFFFFFFFFu;
                for (int at = 0; at < a.A && acc; ++at) {
                    const uint32_t code = a.wide[at] ? ((const uint16_t*)a.cols[at])[r] : ((const uint8_t*)a.cols[at])[r];
                    acc &= __ldg(mk + (long long)(a.bin_off[at] + code) * W + w);
                }
                if (w == W - 1 && (P & 31)) acc &= (1u << (P & 31)) - 1u;
                if (acc) path = w * 32 + (__ffs(acc) - 1);
            }
            if (path < 0) {
                unmatched = true;
                break;
            }
            const int c = a.path_class[a.path_off[m] + path];
            votes[c] = __dadd_rn(votes[c], a.weights[m]);   // votes.put(predClass, count + weight), in member order
            voted[c] = true;
            last = c;
        }
        uint8_t pred = 0xFF;
        if (!unmatched) {
            if (a.M == 1) {
                pred = (uint8_t)last;
            } else {
                // Collections.sort on (count descending) is stable over votes.keySet() order
                int bestc = -1;
                for (int c = 0; c < a.n_classes; ++c) {
                    if (!voted[c]) continue;
                    if (bestc < 0 || votes[c] > votes[bestc] || (votes[c] == votes[bestc] && a.class_rank[c] < a.class_rank[bestc])) bestc = c;
                }
                pred = (uint8_t)bestc;
            }
        }
        a.out[r] = pred;
    }
}

}  // namespace avt
