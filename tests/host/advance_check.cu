// Host-side check of nfb::advance() (nerfacc_b200/csrc/grid.cuh) against the serial marching
// loop it replaces (reference csrc/grid.cu:157-161, 199-203, 208-262).  Built and run by
// tests/test_advance_host.py with:  nvcc -O2 -Xcompiler -ffp-contract=off advance_check.cu
// Exit code 0 and a line "ok <cases> <steps>" on success.
#include <cmath>
#include <cstring>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>

#define NFB_HOST_TEST 1
#include "../../nerfacc_b200/csrc/grid.cuh"

struct Rec { std::vector<float> edges; };

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint64_t rnd() { rng_state ^= rng_state << 7; rng_state ^= rng_state >> 9; return rng_state * 0x2545F4914F6CDD1Dull; }
static double urand() { return (rnd() >> 11) * (1.0 / 9007199254740992.0); }

static float serial(float t, float dt, float target, int cap, int& steps, std::vector<float>* out) {
    while (steps < cap) {
        if (std::fmaf(dt, 0.5f, t) >= target) break;
        volatile float t1 = t + dt;
        if (!(t1 > t)) break;
        const float t1v = t1;
        if (out) out->push_back(t1v);
        t = t1v; ++steps;
    }
    return t;
}

int main(int argc, char** argv) {
    long cases = argc > 1 ? atol(argv[1]) : 2000000;
    long long total_steps = 0;
    const float dts[] = {5e-3f, 1e-3f, 0.013f, 2e-2f, 0.0078125f, 0.75f, 1e-4f, 3.0517578125e-05f, 0.1f, 1.5f};
    for (long c = 0; c < cases; ++c) {
        float dt;
        const int pick = (int)(rnd() % 14);
        if (pick < 10) dt = dts[pick];
        else dt = (float)(std::exp(urand() * 12.0 - 10.0));
        if (rnd() % 16 == 0) {  // dt with few mantissa bits: provokes exact ties
            uint32_t b; std::memcpy(&b, &dt, 4); b &= 0xFFFF0000u << (rnd() % 8); std::memcpy(&dt, &b, 4);
            if (!(dt > 0)) dt = 0.25f;
        }
        float t;
        const int tp = (int)(rnd() % 8);
        if (tp == 0) t = 0.0f;
        else if (tp == 1) t = (float)(-urand() * 3.0 * dt);
        else if (tp == 2) { t = (float)std::ldexp(1.0, (int)(rnd() % 12) - 6); t = std::nextafterf(t, 0.0f); for (int i = (int)(rnd() % 4); i > 0; --i) t = std::nextafterf(t, 0.0f); }
        else t = (float)(std::exp(urand() * 10.0 - 6.0));
        if (rnd() % 8 == 0 && t > 0.0f) {
            // dt exactly halfway between two points of t's grid: every step in t's binade is a
            // rounding tie (to the even mantissa)
            int e; std::frexp(t, &e);
            const float u = (float)std::ldexp(1.0, e - 24);  // ulp of t's binade
            const double n = (double)(rnd() % (1u << (4 + rnd() % 18)));
            dt = (float)((n + 0.5) * u);
        }
        const double span = (rnd() % 4 == 0) ? urand() * 3000.0 : urand() * 40.0;
        float target = (float)(t + span * dt + (urand() - 0.5) * dt);
        if (rnd() % 64 == 0) target = t;  // already there
        int cap = (rnd() % 8 == 0) ? (int)(rnd() % 50) + 1 : 0x7fffffff;
        int s_ref = 0, s_new = 0;
        std::vector<float> e_ref, e_new;
        const float r_ref = serial(t, dt, target, cap, s_ref, &e_ref);
        const float r_new = nfb::advance(t, dt, target, cap, s_new, [&](float t0, float q, int k) {
            for (int j = 1; j <= k; ++j) e_new.push_back(std::fmaf((float)j, q, t0));
        });
        bool ok = (r_ref == r_new) && (s_ref == s_new) && (e_ref.size() == e_new.size());
        if (ok) for (size_t i = 0; i < e_ref.size(); ++i) if (e_ref[i] != e_new[i]) { ok = false; break; }
        if (!ok) {
            printf("MISMATCH case %ld: t=%.9g dt=%.9g target=%.9g cap=%d -> serial (t=%.9g, n=%d) advance (t=%.9g, n=%d)\n",
                   c, t, dt, target, cap, r_ref, s_ref, r_new, s_new);
            return 1;
        }
        total_steps += s_ref;
        // the variant that stops at t >= target (the DDA's tdist += delta consumed up to a time)
        int s2_ref = 0, s2_new = 0;
        float t2 = t;
        while (s2_ref < cap && !(t2 >= target)) { volatile float t1 = t2 + dt; if (!(t1 > t2)) break; t2 = t1; ++s2_ref; }
        float last = -1.0f, last_ref = -1.0f;
        { float tt = t; for (int j = 0; j < s2_ref; ++j) { last_ref = tt; volatile float t1 = tt + dt; tt = t1; } }
        const float r2 = nfb::advance<false>(t, dt, target, cap, s2_new,
                                             [&](float t0, float q, int k) { last = std::fmaf((float)(k - 1), q, t0); });
        if (!(r2 == t2 && s2_ref == s2_new && (s2_ref == 0 || last == last_ref))) {
            printf("MISMATCH (t >= target) case %ld: t=%.9g dt=%.9g target=%.9g cap=%d -> serial (t=%.9g, n=%d, last=%.9g) "
                   "advance (t=%.9g, n=%d, last=%.9g)\n", c, t, dt, target, cap, t2, s2_ref, last_ref, r2, s2_new, last);
            return 1;
        }
    }
    printf("ok %ld %lld\n", cases, total_steps);
    return 0;
}
