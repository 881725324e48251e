/* TEST INFRASTRUCTURE ONLY - plain-C restatement of the integer-valued steps of the path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load the library
 * built from this file (oracle/_build/libss_oracle.so), and only as the checker.
 *
 * Each function follows the reference line by line, with the Python loops written as C loops:
 *   ss_oracle_round_durations  model/modules.py:137-140   clamp(round(exp(logd) - 1) * control, 0)
 *   ss_oracle_bucketize        model/modules.py:86-91     torch.bucketize(v, bins), right=False
 *   ss_oracle_length_regulate  model/modules.py:174-197 + utils/tools.py:319-337
 *   ss_oracle_mel_len          model/modules.py:143       d_rounded.sum(-1).int()
 *   ss_oracle_wave_int16       utils/model.py:105-113     (wav * 32768).astype(int16), trim
 * Pinned by tests/test_oracle.py against tests/golden (outputs of the reference itself).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

/* torch.round = round half to even = rintf under the default rounding mode.  expf here is libm's
 * (correctly rounded to within 1 ulp); the reference uses ATen's vectorised exp (Sleef), so the two
 * can differ in the last bit - tests compare on inputs that are not within 1e-5 of a .5 boundary. */
void ss_oracle_round_durations(const float* log_d, float control, int64_t n, float* out) {
    for (int64_t i = 0; i < n; ++i) {
        float v = rintf(expf(log_d[i]) - 1.0f) * control;
        out[i] = v < 0.0f ? 0.0f : v; /* clamp(min=0): NaN propagates, -0.0 stays -0.0 like torch */
        if (v != v) out[i] = v;
    }
}

/* index = number of edges strictly below v (edges ascending); NaN compares false => n_edges. */
void ss_oracle_bucketize(const float* v, int64_t n, const float* edges, int n_edges, int64_t* out) {
    for (int64_t i = 0; i < n; ++i) {
        int lo = 0, hi = n_edges;
        while (lo < hi) {
            int mid = lo + (hi - lo) / 2;
            if (!(edges[mid] >= v[i])) lo = mid + 1; else hi = mid;
        }
        out[i] = lo;
    }
}

/* Returns T (the padded length).  If out == NULL only lengths / T are computed.
 * x [B,L,C] f32, d [B,L] f32 -> out [B,T,C] f32 (zero rows beyond mel_len), mel_len [B] i64,
 * mel2phone [B,T] i64 (-1 on padding).  max_len < 0 means "longest utterance". */
int64_t ss_oracle_length_regulate(const float* x, const float* d, int64_t B, int64_t L, int64_t C,
                                  int64_t max_len, float* out, int64_t* mel_len, int64_t* mel2phone) {
    int64_t T = 0;
    for (int64_t b = 0; b < B; ++b) {
        int64_t total = 0;
        for (int64_t i = 0; i < L; ++i) {
            int64_t r = (int64_t)d[b * L + i]; /* int(): truncation toward zero */
            if (r > 0) total += r;
        }
        mel_len[b] = total;
        if (total > T) T = total;
    }
    if (max_len >= 0) T = max_len;
    if (!out) return T;
    for (int64_t b = 0; b < B; ++b) {
        int64_t t = 0;
        for (int64_t i = 0; i < L && t < T; ++i) {
            int64_t r = (int64_t)d[b * L + i];
            for (int64_t k = 0; k < r && t < T; ++k, ++t) {
                memcpy(out + (b * T + t) * C, x + (b * L + i) * C, sizeof(float) * (size_t)C);
                if (mel2phone) mel2phone[b * T + t] = i;
            }
        }
        for (; t < T; ++t) {
            memset(out + (b * T + t) * C, 0, sizeof(float) * (size_t)C);
            if (mel2phone) mel2phone[b * T + t] = -1;
        }
    }
    return T;
}

/* float32 running sum in index order, then truncation to int32 (torch sums a short row serially
 * only up to its vector width; tests use integer-valued durations where the order cannot matter). */
void ss_oracle_mel_len(const float* d, int64_t B, int64_t L, int32_t* out) {
    for (int64_t b = 0; b < B; ++b) {
        float s = 0.0f;
        for (int64_t i = 0; i < L; ++i) s += d[b * L + i];
        out[b] = (int32_t)s;
    }
}

void ss_oracle_wave_int16(const float* w, int64_t n, float scale, int16_t* out) {
    for (int64_t i = 0; i < n; ++i) {
        float v = w[i] * scale;
        out[i] = (int16_t)(int32_t)v; /* numpy float->int16 on x86: cvttss2si then low 16 bits */
    }
}
