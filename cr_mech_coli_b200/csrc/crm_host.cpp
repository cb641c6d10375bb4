// crm_host.cpp — host-only pieces of the boundary: save schedule, colour bijection,
// ChaCha8-exact agent placement.  No device code.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <utility>
#include <vector>

#include "../../include/crm_b200.h"

namespace crm {

// Save schedule of run_simulation_with_agents (reference src/simulation.rs:239-249) fed to
// cellular_raza's FixedStepsize::from_partial_save_points: every save time maps to the
// iteration round((t - t0)/dt), all arithmetic in f32.
int build_save_schedule(const crm_config* c, std::vector<uint64_t>& out, uint64_t& max_iter, std::string& err) {
    const float t0 = c->t0, dt = c->dt, t_max = c->t_max;
    if (!(dt > 0.0f) || !(t_max >= t0) || !std::isfinite(t_max) || !std::isfinite(t0)) {
        err = "TimeError: invalid time configuration (need dt > 0 and t_max >= t0)";
        return CRM_ERR_TIME;
    }
    std::vector<float> save{t0};
    if (c->n_saves > 0) {
        const float dtsave = (t_max - t0) / (float)(c->n_saves + 1);
        for (uint32_t n = 1; n < c->n_saves + 1; ++n) save.push_back(t0 + (float)n * dtsave);
    }
    save.push_back(t_max);
    out.clear();
    for (float t : save) out.push_back((uint64_t)std::round((t - t0) / dt));
    std::sort(out.begin(), out.end());
    out.erase(std::unique(out.begin(), out.end()), out.end());
    max_iter = out.back();
    return 0;
}

// ---- rand_chacha 0.9 ChaCha8Rng + rand 0.9 sampling (reference Cargo.lock:1739-1810) ----
struct ChaCha8 {
    uint32_t key[8];
    uint64_t counter = 0;
    uint32_t buf[16];
    int idx = 16;

    explicit ChaCha8(uint64_t state) {  // SeedableRng::seed_from_u64: PCG32 fills the key
        for (int i = 0; i < 8; ++i) {
            state = state * 6364136223846793005ULL + 11634580027462260723ULL;
            uint32_t xs = (uint32_t)(((state >> 18) ^ state) >> 27);
            uint32_t rot = (uint32_t)(state >> 59);
            key[i] = (xs >> rot) | (xs << ((32 - rot) & 31));
        }
    }
    static inline uint32_t rotl(uint32_t x, int n) { return (x << n) | (x >> (32 - n)); }
    void block() {
        uint32_t init[16] = {0x61707865u, 0x3320646Eu, 0x79622D32u, 0x6B206574u};
        for (int i = 0; i < 8; ++i) init[4 + i] = key[i];
        init[12] = (uint32_t)counter;
        init[13] = (uint32_t)(counter >> 32);
        init[14] = init[15] = 0;
        uint32_t x[16];
        memcpy(x, init, sizeof(x));
#define QR(a, b, c, d)                                   \
    x[a] += x[b]; x[d] = rotl(x[d] ^ x[a], 16);          \
    x[c] += x[d]; x[b] = rotl(x[b] ^ x[c], 12);          \
    x[a] += x[b]; x[d] = rotl(x[d] ^ x[a], 8);           \
    x[c] += x[d]; x[b] = rotl(x[b] ^ x[c], 7);
        for (int r = 0; r < 4; ++r) {
            QR(0, 4, 8, 12) QR(1, 5, 9, 13) QR(2, 6, 10, 14) QR(3, 7, 11, 15)
            QR(0, 5, 10, 15) QR(1, 6, 11, 12) QR(2, 7, 8, 13) QR(3, 4, 9, 14)
        }
#undef QR
        for (int i = 0; i < 16; ++i) buf[i] = x[i] + init[i];
        counter++;
        idx = 0;
    }
    uint32_t next_u32() {
        if (idx >= 16) block();
        return buf[idx++];
    }
    // UniformFloat<f32>::sample_single
    float range_f32(float low, float high) {
        float scale = high - low;
        for (;;) {
            uint32_t bits = (next_u32() >> 9) | 0x3F800000u;
            float v12;
            memcpy(&v12, &bits, 4);
            volatile float v01 = v12 - 1.0f;
            volatile float prod = v01 * scale;  // separate mul and add, as in Rust
            float res = prod + low;
            if (res < high) return res;
            uint32_t sb;
            memcpy(&sb, &scale, 4);
            sb -= 1;
            memcpy(&scale, &sb, 4);
        }
    }
    // UniformInt<u32>::sample_single_inclusive(0, n-1): Canon's method
    uint32_t below(uint32_t n) {
        if (n == 0) return next_u32();
        uint64_t m = (uint64_t)next_u32() * n;
        uint32_t result = (uint32_t)(m >> 32), lo = (uint32_t)m;
        if (lo > (uint32_t)(0u - n)) {
            uint32_t new_hi = (uint32_t)(((uint64_t)next_u32() * n) >> 32);
            if ((uint64_t)lo + new_hi > 0xFFFFFFFFull) result += 1;
        }
        return result;
    }
};

}  // namespace crm

extern "C" {

void crm_counter_to_color(uint32_t counter, uint8_t rgb[3]) {
    const uint64_t M = 251ull * 251ull * 251ull;
    uint64_t c = ((uint64_t)counter * 157ull * 163ull * 173ull) % M;
    rgb[0] = (uint8_t)(c / (251ull * 251ull));
    c %= 251ull * 251ull;
    rgb[1] = (uint8_t)(c / 251ull);
    rgb[2] = (uint8_t)(c % 251ull);
}

uint32_t crm_color_to_counter(const uint8_t rgb[3]) {
    const uint64_t M = 251ull * 251ull * 251ull;
    uint64_t c = (uint64_t)rgb[0] * 251ull * 251ull + (uint64_t)rgb[1] * 251ull + rgb[2];
    c = (c * 12590168ull) % M;
    c = (c * 13775961ull) % M;
    c = (c * 12157008ull) % M;
    return (uint32_t)c;
}

int crm_generate_positions(uint32_t n_agents, float spring_length, const crm_config* config, uint64_t rng_seed,
                           const float dx[2], float randomize_positions, uint32_t n_vertices, float* out_pos) {
    using crm::ChaCha8;
    if (!config || !out_pos || n_vertices == 0) return CRM_ERR_INVALID;
    ChaCha8 rng(rng_seed);
    const float s = std::min(std::max(randomize_positions, 0.0f), 1.0f);
    const uint32_t sides = (uint32_t)std::ceil(std::sqrt((float)n_agents));
    const float dchunk1 = (config->domain_size[0] - 2.0f * dx[0]) / (float)sides;
    const float dchunk2 = (config->domain_size[1] - 2.0f * dx[1]) / (float)sides;
    // itertools::iproduct!(0..sides, 0..sides).choose_multiple(&mut rng, n_agents)
    std::vector<std::pair<uint32_t, uint32_t>> picked;
    picked.reserve(n_agents);
    uint64_t seen = 0;
    for (uint32_t i = 0; i < sides; ++i)
        for (uint32_t j = 0; j < sides; ++j) {
            if (picked.size() < n_agents) {
                picked.emplace_back(i, j);
            } else {
                uint32_t k = rng.below((uint32_t)(seen - n_agents + 1 + n_agents));
                if (k < picked.size()) picked[k] = {i, j};
            }
            seen++;
        }
    const float half = (float)n_vertices * spring_length / 2.0f;
    const float H = config->domain_height;
    auto s_gen = [&]() -> float { return s == 0.0f ? 1.0f : rng.range_f32(1.0f - s, 1.0f + s); };
    size_t rod = 0;
    for (auto& ix : picked) {
        const float xlow = dx[0] + (float)ix.first * dchunk1;
        const float ylow = dx[1] + (float)ix.second * dchunk2;
        const float lo_x = xlow + half, hi_x = xlow + dchunk1 - half;
        const float lo_y = ylow + half, hi_y = ylow + dchunk2 - half;
        if (!(lo_x < hi_x) || !(lo_y < hi_y)) return CRM_ERR_INVALID;  // rand panics on an empty range
        const float mx = rng.range_f32(lo_x, hi_x);
        const float my = rng.range_f32(lo_y, hi_y);
        const float mz = rng.range_f32(0.4f * H, 0.6f * H);
        const float angle = rng.range_f32(0.0f, 2.0f * 3.14159265358979323846f);
        volatile float hc = half * cosf(angle), hs = half * sinf(angle), hz = half * 0.0f;
        const float p1[3] = {mx - hc, my - hs, mz - hz};
        float* P = out_pos + rod * n_vertices * 3;
        for (int c = 0; c < 3; ++c)  // nalgebra from_fn fills column-major
            for (uint32_t r = 0; r < n_vertices; ++r) {
                volatile float a = (float)r * spring_length;
                a = a * s_gen();
                float b;
                if (c == 0) {
                    volatile float arg = angle * s_gen();
                    b = cosf(arg);
                } else if (c == 1) {
                    volatile float arg = angle * s_gen();
                    b = sinf(arg);
                } else {
                    b = 0.0f;
                }
                volatile float ab = a * b;
                P[r * 3 + c] = p1[c] + ab;
            }
        rod++;
    }
    return 0;
}

}  // extern "C"

// ---- recycled host blocks (crm_hvec's allocator) ------------------------------------------------------------------
// Blocks of 1 MB and more are kept when freed (at most HOST_CACHE_LIMIT bytes, best fit within 2x on reuse): their
// pages are resident, so the next simulation's mirrors do not fault them in again.  Smaller blocks go straight to
// operator new / delete.
namespace {
constexpr size_t HOST_CACHE_MIN = (size_t)1 << 20, HOST_CACHE_LIMIT = (size_t)1 << 30;
struct HostBlockCache {
    std::mutex m;
    std::vector<std::pair<void*, size_t>> free_blocks;   // (block, capacity)
    std::vector<std::pair<void*, size_t>> live_big;      // capacity of the big blocks handed out
    size_t held = 0;
    ~HostBlockCache() {
        for (auto& b : free_blocks) ::operator delete(b.first);
    }
};
HostBlockCache& host_cache() {
    static HostBlockCache* c = new HostBlockCache();  // (never destroyed: vectors of static objects may outlive it)
    return *c;
}
}  // namespace

void* crm_host_block_alloc(size_t bytes) {
    if (bytes < HOST_CACHE_MIN) return ::operator new(bytes ? bytes : 1);
    HostBlockCache& c = host_cache();
    {
        std::lock_guard<std::mutex> lock(c.m);
        size_t best = SIZE_MAX, best_i = 0;
        for (size_t i = 0; i < c.free_blocks.size(); ++i) {
            const size_t cap = c.free_blocks[i].second;
            if (cap >= bytes && cap <= 2 * bytes && cap < best) { best = cap; best_i = i; }
        }
        if (best != SIZE_MAX) {
            void* p = c.free_blocks[best_i].first;
            c.free_blocks.erase(c.free_blocks.begin() + (long)best_i);
            c.held -= best;
            c.live_big.emplace_back(p, best);
            return p;
        }
    }
    void* p = ::operator new(bytes);
    std::lock_guard<std::mutex> lock(c.m);
    c.live_big.emplace_back(p, bytes);
    return p;
}

void crm_host_block_free(void* p, size_t bytes) noexcept {
    if (!p) return;
    if (bytes < HOST_CACHE_MIN) {
        ::operator delete(p);
        return;
    }
    HostBlockCache& c = host_cache();
    size_t cap = 0;
    {
        std::lock_guard<std::mutex> lock(c.m);
        for (size_t i = 0; i < c.live_big.size(); ++i)
            if (c.live_big[i].first == p) {
                cap = c.live_big[i].second;
                c.live_big.erase(c.live_big.begin() + (long)i);
                break;
            }
        if (cap && c.held + cap <= HOST_CACHE_LIMIT) {
            c.free_blocks.emplace_back(p, cap);
            c.held += cap;
            return;
        }
    }
    ::operator delete(p);
}
