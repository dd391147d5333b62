/*
 * TEST INFRASTRUCTURE -- not part of the product. Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may build, load or call this file.
 *
 * Plain-C restatement of the reference's batch assembly for the training hot path
 * (paths relative to the reference root, FT.py = finetuner-workflow/finetuner/finetuner.py):
 *   - TokenizedDataset.load            FT.py:662-692  slice one context, widen uint16 -> int64, mask rule
 *   - collector                        FT.py:1030-1035 stack ids / mask / labels (labels = a second copy of ids)
 *   - ModifiedTrainer.compute_loss     FT.py:476-477  labels[~attention_mask] = -100
 *   - HF causal shift                  transformers loss_utils.ForCausalLMLoss: target[t] = labels[t+1], last = -100
 * Parity pin: checked against fixtures produced by executing the reference's own class/function bodies
 * (tests/golden/make_golden.py lifts them from the reference file by AST).
 */
#include <stdint.h>
#include <stddef.h>

/* FT.py:662-692 for one context, then FT.py:1030-1035 + FT.py:476-477 for the batch row. */
void ftk_oracle_gather(const uint16_t* tokens, int64_t n_ctx, int32_t T, const int64_t* ctx_idx, int32_t B,
                       int32_t pad_id, int32_t pad_is_eos, int32_t shift_labels, int64_t* input_ids, int64_t* labels,
                       uint8_t* mask, int32_t* n_valid) {
  for (int32_t b = 0; b < B; ++b) {
    const int64_t idx = ctx_idx[b];
    const uint16_t* src = tokens + idx * (int64_t)T;       /* FT.py:663-665: bytes [idx*2T, (idx+1)*2T) */
    int64_t* ids = input_ids + (int64_t)b * T;
    int64_t* lab = labels + (int64_t)b * T;
    uint8_t* msk = mask + (int64_t)b * T;
    /* FT.py:674: ambiguous padding -> only the final context is assumed padded */
    const int all_ones = pad_is_eos && (idx != n_ctx - 1);
    int32_t cnt = 0;
    for (int32_t t = 0; t < T; ++t) {
      ids[t] = (int64_t)src[t];                            /* FT.py:666-672: ushort -> int_ */
      msk[t] = (uint8_t)(all_ones || (int32_t)src[t] != pad_id); /* FT.py:688-691 */
    }
    for (int32_t t = 0; t < T; ++t) {
      if (!shift_labels) {
        lab[t] = msk[t] ? ids[t] : -100;                   /* FT.py:1034 then FT.py:476-477 */
      } else {
        lab[t] = (t + 1 < T && msk[t + 1]) ? ids[t + 1] : -100; /* HF shift of the masked labels */
      }
      cnt += (lab[t] != -100);
    }
    if (n_valid) n_valid[b] = cnt;
  }
}
