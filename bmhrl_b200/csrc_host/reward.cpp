// libbmhrl_reward.so -- host-side reward scorer of the BMHRL RL step (include/bmhrl_reward.h).
//
// The reference recomputes, for every prefix of every sampled caption, the full CIDEr score from Python string
// n-gram dictionaries (metrics/cider.py:28-70, :153-239; metrics/util.py:16-33): O(L^2) dict work per caption,
// 38 % of its RL step.  Here words are int32 ids, an n-gram (n <= 4) is one 128-bit key, the document-frequency
// table is a flat open-addressing table built once, and appending a word to a hypothesis updates the tf-idf norm
// and the clipped dot product of each n-gram order in O(1): one caption costs O(L * n).
//
// Arithmetic follows the reference's double / float32 operations one by one (pow(M_E, x) for the length penalty,
// pow(v, 2) for the norms, fp32 accumulation in the discount scans); compile with -ffp-contract=off.
#include "../../include/bmhrl_reward.h"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace {

thread_local std::string g_err;
int fail(int rc, const std::string& msg) { g_err = msg; return rc; }

constexpr int kMaxN = 4;
constexpr uint32_t kNoWord = 0xFFFFFFFFu;

struct Key {
    uint32_t w[kMaxN];
    bool operator==(const Key& o) const { return std::memcmp(w, o.w, sizeof(w)) == 0; }
};

inline Key make_key(const int* words, int len) {
    Key k;
    for (int i = 0; i < kMaxN; ++i) k.w[i] = i < len ? (uint32_t)words[i] : kNoWord;
    return k;
}

inline uint64_t mix64(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}
inline uint64_t hash_key(const Key& k) {
    uint64_t a = ((uint64_t)k.w[0] << 32) | k.w[1], b = ((uint64_t)k.w[2] << 32) | k.w[3];
    return mix64(a ^ mix64(b + 0x9e3779b97f4a7c15ULL));
}

// Open-addressing map Key -> index into an insertion-ordered entry vector (dict order matters for the sums).
template <class V>
struct OrderedMap {
    std::vector<int32_t> slot;   // -1 = empty, else index into keys/vals
    std::vector<Key> keys;
    std::vector<V> vals;
    uint64_t mask = 0;

    void reserve_pow2(size_t n_entries) {
        size_t cap = 16;
        while (cap < 2 * n_entries + 2) cap <<= 1;
        if (cap != slot.size()) slot.assign(cap, -1); else std::fill(slot.begin(), slot.end(), -1);
        mask = cap - 1;
        keys.clear(); vals.clear();
    }
    void grow() {
        size_t cap = slot.size() * 2;
        slot.assign(cap, -1);
        mask = cap - 1;
        for (size_t i = 0; i < keys.size(); ++i) {
            uint64_t h = hash_key(keys[i]) & mask;
            while (slot[h] >= 0) h = (h + 1) & mask;
            slot[h] = (int32_t)i;
        }
    }
    int32_t find(const Key& k) const {
        if (slot.empty()) return -1;
        uint64_t h = hash_key(k) & mask;
        while (slot[h] >= 0) {
            if (keys[slot[h]] == k) return slot[h];
            h = (h + 1) & mask;
        }
        return -1;
    }
    // returns the entry index, inserting V{} if absent
    int32_t find_or_insert(const Key& k) {
        if (slot.empty()) reserve_pow2(8);
        if (2 * (keys.size() + 1) > slot.size()) grow();
        uint64_t h = hash_key(k) & mask;
        while (slot[h] >= 0) {
            if (keys[slot[h]] == k) return slot[h];
            h = (h + 1) & mask;
        }
        slot[h] = (int32_t)keys.size();
        keys.push_back(k);
        vals.push_back(V{});
        return slot[h];
    }
};

struct DfEntry { long long count; double log_df; };

struct Cider {
    int n;
    double sigma;
    OrderedMap<DfEntry> df;   // n-grams seen more than once in the corpus (metrics/cider.py:112-120)
};

// np.log(max(1.0, document_frequency[ngram])) of metrics/cider.py:167 (0 for n-grams outside the table)
inline double idf_of(const Cider& c, const Key& k) {
    int32_t i = c.df.find(k);
    return i < 0 ? 0.0 : c.df.vals[i].log_df;
}

struct RefEntry { int tf; double v; };            // v = tf * (ref_len - df), ref_len = log(1) = 0
struct HypEntry { int tf; double idf; double vr; };  // vr = weight of the same n-gram in the reference (0 if absent)

// Per-thread scratch for one caption row.
struct RowScratch {
    OrderedMap<RefEntry> ref[kMaxN];
    OrderedMap<HypEntry> hyp[kMaxN];
    std::vector<int> hw;   // hypothesis words so far
};

// One (hypothesis, reference) row of CiderScorer._cider_diff (metrics/cider.py:33-64).
void score_row(const Cider& c, const long long* pred, int L, const int* tok_off, const int* tok_words,
               const unsigned char* tok_is_end, int n_tok, const int* rw, int R, double* rewards, RowScratch& s) {
    const int N = c.n;
    double norm_ref[kMaxN];
    // counts2vec(ref) :153-179 -- dict order = first occurrence, n-gram order by order
    for (int n = 0; n < N; ++n) {
        auto& m = s.ref[n];
        m.reserve_pow2((size_t)std::max(R, 1));
        for (int i = 0; i + n + 1 <= R; ++i) m.vals[m.find_or_insert(make_key(rw + i, n + 1))].tf += 1;
        double nrm = 0.0;
        for (size_t e = 0; e < m.keys.size(); ++e) {
            double dfv = idf_of(c, m.keys[e]);
            m.vals[e].v = (double)m.vals[e].tf * (0.0 - dfv);
            nrm += std::pow(m.vals[e].v, 2.0);
        }
        norm_ref[n] = std::sqrt(nrm);
        s.hyp[n].reserve_pow2((size_t)std::max(L, 1));
    }
    const double length_ref = N >= 2 ? (double)std::max(R - 1, 0) : 0.0;  // `if n == 1: length += term_freq` :176
    const double two_sigma2 = 2.0 * std::pow(c.sigma, 2.0);

    double norm2_h[kMaxN] = {0, 0, 0, 0}, dot[kMaxN] = {0, 0, 0, 0};
    s.hw.clear();
    int n_scored = 0;
    double last = 0.0;
    for (int l = 0; l < L; ++l) {
        long long t = pred[l];
        if (t < 0 || t >= n_tok) t = 0;  // validated by the caller; defensive
        if (tok_is_end[t]) {             // :44-47
            if (n_scored == 0) { rewards[0] = (double)(float)(-0.1); n_scored = 1; last = rewards[0]; }
            break;
        }
        for (int wi = tok_off[t]; wi < tok_off[t + 1]; ++wi) {
            s.hw.push_back(tok_words[wi]);
            const int m = (int)s.hw.size();
            for (int n = 0; n < N && n + 1 <= m; ++n) {  // the one new n-gram of each order ends at the new word
                Key k = make_key(s.hw.data() + (m - n - 1), n + 1);
                auto& hm = s.hyp[n];
                size_t before = hm.keys.size();
                int32_t e = hm.find_or_insert(k);
                HypEntry& h = hm.vals[e];
                if (hm.keys.size() != before) {
                    h.tf = 0;
                    h.idf = idf_of(c, k);
                    int32_t r = s.ref[n].find(k);
                    h.vr = r < 0 ? 0.0 : s.ref[n].vals[r].v;
                }
                if (h.idf == 0.0) { h.tf += 1; continue; }  // weight 0: contributes nothing to norm or dot product
                double v_old = (double)h.tf * (0.0 - h.idf);
                h.tf += 1;
                double v_new = (double)h.tf * (0.0 - h.idf);
                norm2_h[n] += std::pow(v_new, 2.0) - std::pow(v_old, 2.0);
                dot[n] += std::min(v_new, h.vr) * h.vr - (h.tf > 1 ? std::min(v_old, h.vr) * h.vr : 0.0);
            }
        }
        // sim() :181-207 + mean over the n-gram orders :222-226
        const int m = (int)s.hw.size();
        const double length_h = N >= 2 ? (double)std::max(m - 1, 0) : 0.0;
        const double delta = length_h - length_ref;
        const double penalty = std::pow(M_E, -(std::pow(delta, 2.0)) / two_sigma2);
        double sum = 0.0;
        for (int n = 0; n < N; ++n) {
            double val = dot[n];
            double nh = std::sqrt(norm2_h[n]);
            if (nh != 0.0 && norm_ref[n] != 0.0) val /= (nh * norm_ref[n]);
            val *= penalty;
            sum += val;
        }
        last = sum / (double)N;
        rewards[l] = last;
        n_scored = l + 1;
    }
    for (int l = n_scored; l < L; ++l) rewards[l] = last;  // F.pad(..., scores[last_symbol]) :56-59
}

struct BleuScratch {
    OrderedMap<int> ref[kMaxN];   // n-gram -> count in the (single) reference (cook_refs_bleu maxcounts)
    OrderedMap<int> hyp[kMaxN];
    std::vector<int> hw;
};

// One row of BleuScorer._bleu_diff (metrics/bleu.py:26-59): BLEU-1..n of every prefix against ONE reference,
// compute_score (:216-279) with option "average" (one reference), averaged with weights 1/n in fp32 (:276-278).
// Unlike the CIDEr variant the prefix loop does not stop at "</s>".
void bleu_row(int N, const long long* pred, int L, const int* tok_off, const int* tok_words, const int* rw, int R,
              float* rewards, BleuScratch& s) {
    for (int n = 0; n < N; ++n) {
        s.ref[n].reserve_pow2((size_t)std::max(R, 1));
        for (int i = 0; i + n + 1 <= R; ++i) s.ref[n].vals[s.ref[n].find_or_insert(make_key(rw + i, n + 1))] += 1;
        s.hyp[n].reserve_pow2((size_t)std::max(L, 1));
    }
    const double small = 1e-9, tiny = 1e-15;
    const double reflen = (double)R / 1.0;  // float(sum(reflens)) / len(reflens)
    long long correct[kMaxN] = {0, 0, 0, 0};
    s.hw.clear();
    for (int l = 0; l < L; ++l) {
        long long t = pred[l];
        for (int wi = tok_off[t]; wi < tok_off[t + 1]; ++wi) {
            s.hw.push_back(tok_words[wi]);
            const int m = (int)s.hw.size();
            for (int n = 0; n < N && n + 1 <= m; ++n) {
                Key k = make_key(s.hw.data() + (m - n - 1), n + 1);
                int32_t r = s.ref[n].find(k);
                int c = (s.hyp[n].vals[s.hyp[n].find_or_insert(k)] += 1);
                if (r >= 0 && c <= s.ref[n].vals[r]) correct[n] += 1;  // sum of min(refmax, count) grows by one
            }
        }
        const int testlen = (int)s.hw.size();
        double bleu = 1.0, bleus[kMaxN];
        for (int k = 0; k < N; ++k) {
            double guess = (double)std::max(0, testlen - k);
            bleu *= ((double)correct[k] + tiny) / (guess + small);
            bleus[k] = std::pow(bleu, 1.0 / (double)(k + 1));
        }
        const double ratio = ((double)testlen + tiny) / (reflen + small);
        if (ratio < 1.0) {
            const double bp = std::exp(1.0 - 1.0 / ratio);
            for (int k = 0; k < N; ++k) bleus[k] *= bp;
        }
        const float w = (float)(1.0 / (double)N);
        float acc = 0.0f;
        for (int k = 0; k < N; ++k) acc = acc + (float)bleus[k] * w;
        rewards[l] = acc;
    }
}

}  // namespace

extern "C" {

int bmhrl_reward_abi_version(void) { return 1; }
const char* bmhrl_reward_last_error(void) { return g_err.c_str(); }

int bmhrl_cider_create(const int* corpus_words, const long long* corpus_off, long long n_caps, int n, double sigma,
                       void** out) {
    if (!out) return fail(-1, "bmhrl_cider_create: out is null");
    if (n < 1 || n > kMaxN) return fail(-1, "bmhrl_cider_create: n must be in 1..4");
    if (n_caps < 0 || (n_caps > 0 && (!corpus_words || !corpus_off))) return fail(-1, "bmhrl_cider_create: bad corpus");
    try {
        OrderedMap<long long> counts;
        long long total = n_caps > 0 ? corpus_off[n_caps] - corpus_off[0] : 0;
        counts.reserve_pow2((size_t)std::max<long long>(total, 8));
        for (long long c = 0; c < n_caps; ++c) {
            const int* w = corpus_words + corpus_off[c];
            long long len = corpus_off[c + 1] - corpus_off[c];
            if (len < 0) return fail(-1, "bmhrl_cider_create: corpus_off not monotone");
            for (int k = 1; k <= n; ++k)
                for (long long i = 0; i + k <= len; ++i) {
                    for (int j = 0; j < k; ++j)
                        if (w[i + j] < 0) return fail(-1, "bmhrl_cider_create: negative word id");
                    counts.vals[counts.find_or_insert(make_key(w + i, k))] += 1;
                }
        }
        Cider* c = new Cider();
        c->n = n;
        c->sigma = sigma;
        size_t kept = 0;
        for (size_t i = 0; i < counts.keys.size(); ++i) kept += counts.vals[i] > 1;
        c->df.reserve_pow2(std::max<size_t>(kept, 8));
        for (size_t i = 0; i < counts.keys.size(); ++i)
            if (counts.vals[i] > 1) {
                DfEntry& e = c->df.vals[c->df.find_or_insert(counts.keys[i])];
                e.count = counts.vals[i];
                e.log_df = std::log(std::max(1.0, (double)counts.vals[i]));
            }
        *out = c;
        return 0;
    } catch (const std::exception& e) {
        return fail(-3, std::string("bmhrl_cider_create: ") + e.what());
    }
}

int bmhrl_cider_destroy(void* scorer) {
    delete static_cast<Cider*>(scorer);
    return 0;
}

long long bmhrl_cider_df_size(void* scorer) {
    return scorer ? (long long)static_cast<Cider*>(scorer)->df.keys.size() : -1;
}

long long bmhrl_cider_df_lookup(void* scorer, const int* ngram, int len) {
    if (!scorer || !ngram || len < 1 || len > kMaxN) return -1;
    const Cider& c = *static_cast<Cider*>(scorer);
    int32_t i = c.df.find(make_key(ngram, len));
    return i < 0 ? 0 : c.df.vals[i].count;
}

int bmhrl_cider_prefix_rewards(void* scorer, const long long* pred, int B, int L, const int* tok_off,
                               const int* tok_words, const unsigned char* tok_is_end, int n_tok, const int* ref_words,
                               const long long* ref_off, double* rewards, float* delta, int n_threads) {
    if (!scorer) return fail(-1, "bmhrl_cider_prefix_rewards: null scorer");
    if (B < 0 || L < 1) return fail(-1, "bmhrl_cider_prefix_rewards: need B >= 0, L >= 1");
    if (B == 0) return 0;
    if (!pred || !tok_off || !tok_is_end || !ref_off || !rewards || !delta)
        return fail(-1, "bmhrl_cider_prefix_rewards: null argument");
    for (long long i = 0; i < (long long)B * L; ++i)
        if (pred[i] < 0 || pred[i] >= n_tok) return fail(-2, "bmhrl_cider_prefix_rewards: token id outside the vocabulary");
    const Cider& c = *static_cast<Cider*>(scorer);
    try {
        int hw = (int)std::thread::hardware_concurrency();
        int T = n_threads > 0 ? n_threads : std::max(1, std::min(hw > 0 ? hw : 1, B / 8));
        T = std::min(T, B);
        auto work = [&](int b0, int b1) {
            RowScratch s;
            for (int b = b0; b < b1; ++b) {
                int R = (int)(ref_off[b + 1] - ref_off[b]);
                double* rw = rewards + (size_t)b * L;
                score_row(c, pred + (size_t)b * L, L, tok_off, tok_words, tok_is_end, n_tok, ref_words + ref_off[b], R,
                          rw, s);
                float* d = delta + (size_t)b * L;
                d[0] = (float)rw[0];
                for (int l = 1; l < L; ++l) d[l] = (float)(rw[l] - rw[l - 1]);  // :66-67, fp64 difference then .float()
            }
        };
        if (T <= 1) {
            work(0, B);
        } else {
            std::vector<std::thread> th;
            for (int t = 0; t < T; ++t) th.emplace_back(work, (int)((long long)B * t / T), (int)((long long)B * (t + 1) / T));
            for (auto& x : th) x.join();
        }
        return 0;
    } catch (const std::exception& e) {
        return fail(-3, std::string("bmhrl_cider_prefix_rewards: ") + e.what());
    }
}

int bmhrl_bleu_prefix_rewards(const long long* pred, int B, int L, const int* tok_off, const int* tok_words, int n_tok,
                              const int* ref_words, const long long* ref_off, int n, float* rewards, float* delta,
                              int n_threads) {
    if (n < 1 || n > kMaxN) return fail(-1, "bmhrl_bleu_prefix_rewards: n must be in 1..4");
    if (B < 0 || L < 1) return fail(-1, "bmhrl_bleu_prefix_rewards: need B >= 0, L >= 1");
    if (B == 0) return 0;
    if (!pred || !tok_off || !ref_off || !rewards || !delta) return fail(-1, "bmhrl_bleu_prefix_rewards: null argument");
    for (long long i = 0; i < (long long)B * L; ++i)
        if (pred[i] < 0 || pred[i] >= n_tok) return fail(-2, "bmhrl_bleu_prefix_rewards: token id outside the vocabulary");
    try {
        int hw = (int)std::thread::hardware_concurrency();
        int T = n_threads > 0 ? n_threads : std::max(1, std::min(hw > 0 ? hw : 1, B / 8));
        T = std::min(T, B);
        auto work = [&](int b0, int b1) {
            BleuScratch s;
            for (int b = b0; b < b1; ++b) {
                float* rw = rewards + (size_t)b * L;
                bleu_row(n, pred + (size_t)b * L, L, tok_off, tok_words, ref_words + ref_off[b],
                         (int)(ref_off[b + 1] - ref_off[b]), rw, s);
                float* d = delta + (size_t)b * L;
                d[0] = rw[0];
                for (int l = 1; l < L; ++l) d[l] = rw[l] - rw[l - 1];  // fp32 tensors in the reference (:52-53)
            }
        };
        if (T <= 1) {
            work(0, B);
        } else {
            std::vector<std::thread> th;
            for (int t = 0; t < T; ++t) th.emplace_back(work, (int)((long long)B * t / T), (int)((long long)B * (t + 1) / T));
            for (auto& x : th) x.join();
        }
        return 0;
    } catch (const std::exception& e) {
        return fail(-3, std::string("bmhrl_bleu_prefix_rewards: ") + e.what());
    }
}

int bmhrl_discount_rewards(const float* diff, int B, int L, double gamma, int n_step, float* out) {
    if (B < 0 || L < 0 || (B > 0 && L > 0 && (!diff || !out))) return fail(-1, "bmhrl_discount_rewards: bad argument");
    int K = std::min(L, std::max(n_step, 0));
    std::vector<float> g((size_t)std::max(K, 1));
    for (int i = 0; i < K; ++i) g[i] = (float)std::pow(gamma, (double)i);  // python float ** int, cast by the fp32 mul
    for (int b = 0; b < B; ++b) {
        const float* r = diff + (size_t)b * L;
        for (int l = 0; l < L; ++l) {
            float acc = 0.0f;
            for (int i = 0; l + i < L && i < n_step; ++i) acc = acc + g[i] * r[l + i];
            out[(size_t)b * L + l] = acc;
        }
    }
    return 0;
}

int bmhrl_segment_reward(const float* reward, const unsigned char* sections, int B, int L, float* out) {
    if (B < 0 || L < 0 || (B > 0 && L > 0 && (!reward || !sections || !out)))
        return fail(-1, "bmhrl_segment_reward: bad argument");
    for (int b = 0; b < B; ++b) {
        const float* r = reward + (size_t)b * L;
        float* o = out + (size_t)b * L;
        int old_l = 0;
        for (int l = 0; l < L; ++l) o[l] = 0.0f;
        for (int l = 0; l < L; ++l)
            if (sections[(size_t)b * L + l]) {
                float sum = 0.0f;
                for (int j = old_l; j <= l; ++j) sum += r[j];
                for (int j = old_l; j <= l; ++j) o[j] = sum;
                old_l = l + 1;
            }
    }
    return 0;
}

int bmhrl_discount_segments(float* diff, const unsigned char* sections, int B, int L, double gamma) {
    if (B < 0 || L < 0 || (B > 0 && L > 0 && (!diff || !sections)))
        return fail(-1, "bmhrl_discount_segments: bad argument");
    // torch.nonzero(segments): row-major list of segment ends
    std::vector<std::pair<int, int>> seg;
    for (int b = 0; b < B; ++b)
        for (int l = 0; l < L; ++l)
            if (sections[(size_t)b * L + l]) seg.emplace_back(b, l);
    int old_l = 0, old_b = 0;
    for (size_t i = 0; i < seg.size(); ++i) {
        int b = seg[i].first, l = seg[i].second;
        float* row = diff + (size_t)b * L;
        if (old_b != b) {
            for (int j = old_l; j < L; ++j) diff[(size_t)old_b * L + j] = 0.0f;
            old_l = 0;
            old_b = b;
        }
        if (i < 2) {  // `if i < segment_idx.shape[0]` -- segment_idx is one (b, l) pair, so this is i < 2
            for (size_t k = i + 1; k < seg.size() && seg[k].first == b; ++k) {
                float g = (float)std::pow(gamma, (double)(k - i));
                row[l] = row[l] + g * row[seg[k].second];  // in place through the view cider_diff[b, l]
            }
        }
        float v = row[l];
        for (int j = old_l; j <= l; ++j) row[j] = v;
        old_l = l + 1;
    }
    return 0;
}

}  // extern "C"
