/* bmhrl_reward.h -- C ABI of libbmhrl_reward.so: the per-prefix CIDEr reward scorer of the BMHRL RL step and
 * its discount / segment scans.  HOST code (C++17, no CUDA, no torch): the
 * reference computes these rewards in Python string / dict code on the host between the sampling and the loss of
 * every RL step (epoch_loops/captioning_bmrl_loops.py:219-230, :287), O(L^2) n-gram recounts per caption.  Here
 * words are int32 ids, n-grams are packed 128-bit keys in open-addressing tables and the per-prefix scores are
 * updated incrementally (O(L) per caption).
 *
 * Conventions: plain pointers + sizes, row-major, no allocation visible to the caller except the scorer handle.
 * Every function returns 0 on success, < 0 on error (bmhrl_reward_last_error() gives the text); nothing throws.
 * Word ids are arbitrary non-negative int32 chosen by the caller (the Python side maps strings to ids).
 */
#ifndef BMHRL_REWARD_H
#define BMHRL_REWARD_H

#ifdef __cplusplus
extern "C" {
#endif

int bmhrl_reward_abi_version(void);
const char* bmhrl_reward_last_error(void);

/* metrics/cider.py:112-120 precook_corpus + :14-26 CiderScorer.__init__.
 * corpus_words / corpus_off: CSR of the tokenised training captions (n_caps rows, corpus_off has n_caps+1 entries).
 * Counts every 1..n-gram over the corpus and keeps those seen more than once (the reference's document-frequency
 * table).  Returns an opaque handle through *out. */
int bmhrl_cider_create(const int* corpus_words, const long long* corpus_off, long long n_caps, int n, double sigma,
                       void** out);
int bmhrl_cider_destroy(void* scorer);
/* number of n-grams kept in the document-frequency table (for tests) */
long long bmhrl_cider_df_size(void* scorer);
/* document-frequency count of one n-gram (0 if absent); len in 1..n */
long long bmhrl_cider_df_lookup(void* scorer, const int* ngram, int len);

/* metrics/cider.py:28-70 CiderScorer._cider_diff (with CiderScorerObj.compute_cider :153-239 for one
 * hypothesis / one reference, so ref_len = log(1) = 0).
 *   pred        (B, L) sampled token ids
 *   tok_off / tok_words  CSR: the words token id t contributes to " ".join(hypo).split() (normally exactly one;
 *               zero for whitespace tokens), n_tok rows
 *   tok_is_end  (n_tok) 1 where vocab.itos[t] == "</s>" (the prefix loop stops there, :44-47)
 *   ref_words / ref_off  CSR of target[b].lower().split() as word ids, B rows
 * Outputs: rewards (B, L) double = the per-prefix CIDEr score padded with the last scored prefix (:56-59);
 *          delta (B, L) float = rewards[:, l] - rewards[:, l-1], delta[:, 0] = rewards[:, 0] (:66-67).
 * n_threads <= 0 uses all hardware threads (rows are independent). */
int bmhrl_cider_prefix_rewards(void* scorer, const long long* pred, int B, int L, const int* tok_off,
                               const int* tok_words, const unsigned char* tok_is_end, int n_tok, const int* ref_words,
                               const long long* ref_off, double* rewards, float* delta, int n_threads);

/* metrics/bleu.py:26-59 BleuScorer._bleu_diff (BleuScorerObj.compute_score :216-279 for one hypothesis / one
 * reference, option "average"): BLEU-1..n of every prefix, brevity penalty exp(1 - 1/ratio), mean over the orders
 * in fp32.  The prefix loop does NOT stop at "</s>".  tok_off / tok_words: CSR of vocab.itos[t].lower().split();
 * ref_words / ref_off: CSR of target[b].lower().split().  rewards, delta: (B, L) float. */
int bmhrl_bleu_prefix_rewards(const long long* pred, int B, int L, const int* tok_off, const int* tok_words, int n_tok,
                              const int* ref_words, const long long* ref_off, int n, float* rewards, float* delta,
                              int n_threads);

/* metrics/util.py:56-66 discontinue_reward(cider_diff, gamma, n_step) without segments:
 * out[b, l] = sum_{i < n_step} fl32(gamma^i) * diff[b, l+i], accumulated in fp32 in the reference's order. */
int bmhrl_discount_rewards(const float* diff, int B, int L, double gamma, int n_step, float* out);

/* metrics/batched_meteor.py:19-36 segment_reward: every position of a segment (positions up to and including a
 * nonzero entry of `sections`) gets the sum of the step rewards of that segment; positions after the last segment
 * end of a row stay 0. */
int bmhrl_segment_reward(const float* reward, const unsigned char* sections, int B, int L, float* out);

/* metrics/util.py:67-88 discontinue_reward(cider_diff, gamma, segments=sections), IN PLACE on diff like the
 * reference, including its quirks: the look-ahead discount is applied only to the first two segment ends of the
 * whole batch (`i < segment_idx.shape[0]` compares with 2), a row's tail is zeroed only when a later row has a
 * segment end, and the last such row keeps its tail. */
int bmhrl_discount_segments(float* diff, const unsigned char* sections, int B, int L, double gamma);

#ifdef __cplusplus
}
#endif
#endif
