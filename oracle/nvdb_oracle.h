/* oracle/nvdb_oracle.h — TEST INFRASTRUCTURE ONLY. See nvdb_oracle.c for the header comment. */
#ifndef NVDB_ORACLE_H
#define NVDB_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* dtype codes = reference include/nvdb/vecbin_format.h:10-14 */
enum { ORC_F32 = 1, ORC_F16 = 2, ORC_I8 = 3 };

float orc_f16_to_f32(uint16_t h);
uint16_t orc_f32_to_f16_rne(float f);

float orc_dot_f32(const float* a, const float* b, uint32_t dim);
float orc_dot_f16(const float* q, const uint16_t* x, uint32_t dim);
float orc_dot_i8(const float* q, const int8_t* x, uint32_t dim, float scale);
float orc_score_row(const void* payload, const float* scales, uint32_t dtype, uint64_t row, uint32_t dim,
                    const float* q);
void orc_score_all(const void* payload, const float* scales, uint32_t dtype, uint64_t n, uint32_t dim,
                   const float* q, float* out_scores);

/* returns number of results (min(k,n)); ids/scores sorted by (score desc, id asc) */
uint32_t orc_flat_search(const void* payload, const float* scales, uint32_t dtype, uint64_t n, uint32_t dim,
                         const float* q, uint32_t k, uint64_t* out_ids, float* out_scores);
uint32_t orc_flat_search_mt(const void* payload, const float* scales, uint32_t dtype, uint64_t n, uint32_t dim,
                            const float* q, uint32_t k, uint64_t* out_ids, float* out_scores);
void orc_flat_search_batch(const void* payload, const float* scales, uint32_t dtype, uint64_t n, uint32_t dim,
                           const float* queries, uint32_t nq, uint32_t k, uint64_t* out_ids, float* out_scores);

void orc_quantize_i8_row(const float* row, uint32_t dim, int8_t* out, float* out_scale);
void orc_quantize_i8(const float* rows, uint64_t n, uint32_t dim, int8_t* out, float* out_scales);
void orc_convert_f16(const float* in, uint64_t count, uint16_t* out);

float orc_l2_cpu(const float* q, const void* payload, const float* scales, uint32_t dtype, uint64_t row,
                 uint32_t dim);
float orc_l2_cudaorder_f16(const float* q, const uint16_t* row, uint32_t dim);
/* mode 0: CPU refine (double accumulate, reference apps/nvdb_ivf_eval.cpp:232-307)
   mode 1: CUDA-kernel arithmetic (reference src/cuda_refine.cu:326-382), f16 base only */
void orc_refine_l2(int mode, const void* payload, const float* scales, uint32_t dtype, uint64_t n, uint32_t dim,
                   const float* queries, const uint32_t* cand, uint32_t Q, uint32_t R, uint32_t K,
                   uint32_t* out_ids, float* out_dist);

/* tensor-core scan limb schemes (restatement of nano-vectordb_b200/csrc/scan_mma.cu, not of the reference) */
void orc_query_limbs_i8(const float* q, uint32_t dim, int8_t* q1, int8_t* q2, float* s1_out, float* s2_out);
float orc_dot_i8_limbs(const int8_t* q1, const int8_t* q2, float s1, float s2, const int8_t* x, uint32_t dim,
                       float scale, int32_t* i1_out, int32_t* i2_out);
void orc_score_all_i8_limbs(const int8_t* payload, const float* scales, uint64_t n, uint32_t dim, const float* q,
                            float* out_scores, int32_t* out_i1, int32_t* out_i2);
void orc_query_limbs_f16(const float* q, uint32_t dim, uint16_t* hi, uint16_t* lo, float* qscale_out);

/* synthetic base generator (ours, not the reference's): integer hash -> exactly reproducible anywhere */
int32_t orc_synth_i16(uint64_t seed, uint64_t global_index);
void orc_synth_rows_f32(uint64_t seed, uint64_t row0, uint64_t nrows, uint32_t dim, float* out);

#ifdef __cplusplus
}
#endif
#endif
