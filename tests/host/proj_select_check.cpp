// Host-side check of whisper_rs_b200/csrc/decode_proj_select.cuh (the exact per-thread code of the projection kernel's
// epilogue) against brute force.  Built and run by tests/test_proj_select_host.py with g++; no GPU involved.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../whisper_rs_b200/csrc/decode_proj_select.cuh"

using namespace wrs;

static int fails = 0;
#define CHECK(c, ...) do { if (!(c)) { if (fails++ < 20) { printf("FAIL %s:%d: ", __FILE__, __LINE__); printf(__VA_ARGS__); printf("\n"); } } } while (0)

static float pred(float f) {                 // the largest float below f (what the kernel uses as "floor")
    if (!(f > -INFINITY)) return f;
    if (f == 0.0f) return -1.4e-45f;
    int b; memcpy(&b, &f, 4); b += f > 0.0f ? -1 : 1; memcpy(&f, &b, 4); return f;
}

template <int KM>
static void check_lists(int K, unsigned seed) {
    srand(seed);
    for (int trial = 0; trial < 400; ++trial) {
        const int n = 1 + rand() % 176;
        const int levels = (trial % 3 == 0) ? 4 : (trial % 3 == 1 ? 40 : 100000);     // few levels: ties by the dozen
        std::vector<float> x(n);
        for (auto& v : x) { v = (float)(rand() % levels) * 0.25f - 3.0f; if (rand() % 11 == 0) v = -INFINITY; }
        // a bound some earlier tile gave (maybe none): values >= bound may enter, ties included
        const float bound = (trial % 4 == 0) ? -INFINITY : (float)(rand() % levels) * 0.25f - 3.0f;
        ProjRow<KM> st;
        st.floor = st.kth = pred(bound);
        for (int e = 0; e < KM; ++e) { st.v[e] = e < KM - K ? INFINITY : -INFINITY; st.id[e] = INT_MAX; }
        const int id0 = 1000;
        for (int i = 0; i < n; ++i)
            if (x[i] > st.kth) proj_insert<KM>(st, x[i], id0 + i);      // (the kernel calls it for candidates only; calling it for all is equivalent)
        // brute force: (value desc, id asc) among finite values >= bound, the first K
        std::vector<std::pair<float, int>> ref;
        for (int i = 0; i < n; ++i) if (x[i] > -INFINITY && x[i] >= bound) ref.push_back({x[i], id0 + i});
        std::stable_sort(ref.begin(), ref.end(), [](const std::pair<float, int>& a, const std::pair<float, int>& b) { return a.first > b.first; });
        for (int e = 0; e < K; ++e) {
            const float gv = st.v[KM - K + e]; const int gi = st.id[KM - K + e];
            if (e < (int)ref.size()) CHECK(gv == ref[e].first && gi == ref[e].second, "KM %d K %d trial %d entry %d: got (%g, %d) want (%g, %d)", KM, K, trial, e, gv, gi, ref[e].first, ref[e].second);
            else CHECK(gv == -INFINITY && gi == INT_MAX, "KM %d K %d trial %d entry %d: expected empty, got (%g, %d)", KM, K, trial, e, gv, gi);
        }
        for (int e = 0; e < KM - K; ++e) CHECK(st.v[e] == INFINITY, "sentinel displaced");
    }
}

int main() {
    // range_bits
    for (int lo = -40; lo <= 40; ++lo)
        for (int hi = -40; hi <= 72; ++hi) {
            unsigned want = 0;
            for (int i = 0; i < 32; ++i) if (lo <= i && i < hi) want |= 1u << i;
            CHECK(range_bits(lo, hi) == want, "range_bits(%d, %d)", lo, hi);
        }
    // kth_of_12
    srand(7);
    for (int trial = 0; trial < 2000; ++trial) {
        float h[12];
        for (auto& v : h) { v = (float)(rand() % (trial % 2 ? 5 : 1000)); if (rand() % 5 == 0) v = -INFINITY; }
        float s[12]; memcpy(s, h, sizeof(h));
        std::sort(s, s + 12, [](float a, float b) { return a > b; });
        for (int K = 1; K <= 12; ++K) CHECK(kth_of_12(h, K) == s[K - 1], "kth_of_12 trial %d K %d", trial, K);
    }
    // tile geometry: the tiles partition [0, ts_split) and [ts_split, V) without straddling
    const int cases[][2] = {{50364, 51865}, {50363, 51864}, {51865, 51865}, {176, 177}, {1, 352}};
    for (auto& c : cases) {
        const int ts = c[0], V = c[1];
        const int nt = (ts + kPN - 1) / kPN, n = nt + (V - ts + kPN - 1) / kPN;
        int covered = 0;
        for (int t = 0; t < n; ++t) {
            const int b = proj_tile_base(t, nt, ts), e = proj_tile_end(t, nt, ts, V);
            CHECK(b == covered && e > b && e - b <= kPN && (e <= ts || b >= ts), "tile %d of (%d, %d): [%d, %d)", t, ts, V, b, e);
            covered = e;
        }
        CHECK(covered == V, "tiles of (%d, %d) end at %d", ts, V, covered);
    }
    // compact rows of a short M tile
    for (int R = 1; R <= 700; ++R)
        for (int mt = 0; mt * kPM < R; ++mt) {
            const int r8 = proj_rows8(R, mt), left = R - mt * kPM;
            CHECK(r8 % 8 == 0 && r8 <= kPM && r8 >= (left < kPM ? left : kPM) && r8 < (left < kPM ? left : kPM) + 8, "proj_rows8(%d, %d) = %d", R, mt, r8);
        }
    // the lists: every list size the kernel instantiates, every K it can hold
    const int kms[] = {2, 4, 6, 8, 12, 16};
    for (int K = 1; K <= 16; ++K) {
        if (K <= 2) check_lists<2>(K, 100 + K);
        if (K <= 4) check_lists<4>(K, 200 + K);
        if (K <= 6) check_lists<6>(K, 300 + K);
        if (K <= 8) check_lists<8>(K, 400 + K);
        if (K <= 12) check_lists<12>(K, 500 + K);
        check_lists<16>(K, 600 + K);
    }
    (void)kms;
    printf(fails ? "proj_select_check: %d failures\n" : "proj_select_check: ok%.0d\n", fails);
    return fails ? 1 : 0;
}
