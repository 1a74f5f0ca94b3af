// tp.h — a model split over several GPUs of one node (tensor parallelism, tp.cu)
#pragma once
#include <stdint.h>

#include <memory>
#include <string>
#include <vector>

#include "../../include/aios_b200.h"

namespace aios {

struct TpModelImpl;

struct TpModel {
    // devices: CUDA ordinals, one shard each (2, 4 or 8 of them; any count that divides the model's kv heads, FF and
    // vocabulary as Model's constructor checks)
    TpModel(const std::vector<int> &devices, const std::string &name, const std::string &path, int n_ctx, int flags);
    ~TpModel();
    int n_ranks() const;
    int n_vocab() const;
    int n_ctx() const;
    int n_past() const;
    void reset();
    void prefill(const int32_t *tokens, int n);                       // token by token through the decode pass
    int decode(const aios_sampling &sp, int max_tokens, int32_t *out);  // after a prefill; returns the tokens written
    void read_logits(float *out);                                     // rank 0's copy of the gathered logits
    void bench_decode(int n_tokens, aios_tp_bench *out);              // greedy steps with the all-reduces timed
    void info(aios_model_info *out) const;  // the unsplit model's shape; weight / KV bytes of ONE shard

   private:
    std::unique_ptr<TpModelImpl> impl;
};

}  // namespace aios
