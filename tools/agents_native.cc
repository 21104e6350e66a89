// "A agents x one frame in flight" through the C-ABI from native host threads - the shape in which the reference feeds
// its front end (src/MultiAgentSystem.cc:56-66: one Tracking thread per agent; src/Agent.cc:92-154: the thread grabs ONE
// image, extracts, tracks against the last frame, grabs the next).  Each agent thread owns an extractor handle, a matcher
// handle and one frame slot on the same GPU and makes the two synchronous calls a Tracking thread makes per frame:
//   orbx_extract            = ORBextractor::operator()                 (src/ORBextractor.cc:1086-1168)
//   orbm_frame_update + orbm_search_frame = Frame grid + SearchByProjection(Cur, Last, th, mono) (src/ORBmatcher.cc:1676-1887)
// with HOST buffers in and out.  bench.py's Python leg measures the same loop under the GIL; this one shows what the
// library and the GPU do when the host side is native, as it is in the reference.
//
// usage: agents_native <frames.u8> <width> <height> <nframes_in_file> <device> <frames_per_agent> <A1,A2,...> [track=1] [pin=0]
// pin = 1 page-locks the frame array (orbx_pin_host: a camera ring buffer registered once), so frames are uploaded in place.
// prints one JSON object.  Built by __graft_entry__.build() (g++, links libmam3slam_b200.so).
#include "../include/orbx.h"
#include "../include/orbm.h"
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

typedef std::chrono::steady_clock Clock;
static double secondsSince(Clock::time_point t0) { return std::chrono::duration<double>(Clock::now() - t0).count(); }

struct FrameOut {
    std::vector<orbx_keypoint> kps;
    std::vector<uint8_t> desc;
    std::vector<float> u, v, ang;
    std::vector<int> oct;
    int n = 0;
};

struct Agent {
    orbx_handle ex = nullptr;
    orbm_handle mt = nullptr;
    orbm_frame slot = nullptr;
    FrameOut cur, last;
    std::vector<uint8_t> occ, valid;
    std::vector<int> kpMatch;
    std::vector<float> scale;
    std::vector<double> lat;
    double tExtract = 0, tUpdate = 0, tSearch = 0;      // seconds inside each call of the timed frames
    long long matches = 0;
    int cap = 0;
    std::string error;
};

static bool fail(Agent& a, const char* what, int rc, const char* msg) {
    a.error = std::string(what) + ": " + std::to_string(rc) + " " + (msg ? msg : "");
    return false;
}

static bool setup(Agent& a, int device) {
    orbx_config cfg = {1000, 1.2f, 8, 20, 7, device, 1};
    int rc = orbx_create(&cfg, &a.ex);
    if (rc) return fail(a, "orbx_create", rc, orbx_strerror(rc));
    if ((rc = orbm_create(device, &a.mt))) return fail(a, "orbm_create", rc, orbm_strerror(rc));
    a.cap = orbx_max_keypoints(a.ex);
    for (FrameOut* f : {&a.cur, &a.last}) {
        f->kps.resize(a.cap); f->desc.resize((size_t)a.cap * 32);
        f->u.resize(a.cap); f->v.resize(a.cap); f->ang.resize(a.cap); f->oct.resize(a.cap);
    }
    a.occ.assign(a.cap, 0); a.valid.assign(a.cap, 1); a.kpMatch.assign(a.cap, -1);
    a.scale.resize(8);
    orbx_get_tables(a.ex, a.scale.data(), nullptr, nullptr, nullptr, nullptr);
    return true;
}

// one frame of an agent's Tracking thread
static bool grab(Agent& a, const uint8_t* img, int w, int h, bool track) {
    int mono = 0;
    Clock::time_point c0 = Clock::now();
    int rc = orbx_extract(a.ex, img, w, h, w, 0, 0, a.cur.kps.data(), a.cur.desc.data(), a.cap, &a.cur.n, &mono);
    if (rc) return fail(a, "orbx_extract", rc, orbx_last_error(a.ex));
    a.tExtract += secondsSince(c0);
    const int n = a.cur.n;
    for (int i = 0; i < n; ++i) {                 // what the projection of the last frame's points reduces to without a pose
        a.cur.u[i] = a.cur.kps[i].x; a.cur.v[i] = a.cur.kps[i].y;
        a.cur.ang[i] = a.cur.kps[i].angle; a.cur.oct[i] = a.cur.kps[i].octave;
    }
    if (track && a.last.n > 0 && n > 0) {
        orbm_frame_view view = {n, a.cur.kps.data(), a.cur.desc.data(), 0.f, 0.f, (float)w, (float)h, 0};
        c0 = Clock::now();
        rc = a.slot ? orbm_frame_update(a.mt, a.slot, &view) : orbm_frame_create(a.mt, &view, &a.slot);
        if (rc) return fail(a, "orbm_frame_update", rc, orbm_last_error(a.mt));
        a.tUpdate += secondsSince(c0);
        std::fill(a.occ.begin(), a.occ.begin() + n, (uint8_t)0);
        int nm = 0;
        c0 = Clock::now();
        rc = orbm_search_frame(a.mt, a.slot, a.occ.data(), a.scale.data(), 8, a.last.n, a.valid.data(), a.last.u.data(),
                               a.last.v.data(), a.last.oct.data(), a.last.ang.data(), a.last.desc.data(), 15.f, 100, 1,
                               a.kpMatch.data(), &nm, nullptr);
        if (rc) return fail(a, "orbm_search_frame", rc, orbm_last_error(a.mt));
        a.tSearch += secondsSince(c0);
        a.matches += nm;
    }
    std::swap(a.cur, a.last);
    return true;
}

static double percentile(std::vector<double> v, double p) {
    if (v.empty()) return 0;
    std::sort(v.begin(), v.end());
    return v[std::min(v.size() - 1, (size_t)(p * (v.size() - 1) + 0.5))];
}

int main(int argc, char** argv) {
    if (argc < 8) { fprintf(stderr, "usage: %s frames.u8 w h nfile device frames_per_agent A1,A2,.. [track]\n", argv[0]); return 2; }
    const int w = atoi(argv[2]), h = atoi(argv[3]), nfile = atoi(argv[4]), device = atoi(argv[5]), per = atoi(argv[6]);
    const bool track = argc > 8 ? atoi(argv[8]) != 0 : true;
    const bool pin = argc > 9 ? atoi(argv[9]) != 0 : false;
    std::vector<int> counts;
    for (char* t = strtok(argv[7], ","); t; t = strtok(nullptr, ",")) counts.push_back(atoi(t));
    std::vector<uint8_t> frames((size_t)w * h * nfile);
    FILE* f = fopen(argv[1], "rb");
    if (!f || fread(frames.data(), 1, frames.size(), f) != frames.size()) { fprintf(stderr, "cannot read %s\n", argv[1]); return 3; }
    fclose(f);
    if (orbx_device_count() <= device) { fprintf(stderr, "no CUDA device %d\n", device); return 4; }
    if (pin && orbx_pin_host(frames.data(), frames.size())) { fprintf(stderr, "orbx_pin_host failed\n"); return 4; }

    printf("{\"track\": %d, \"pinned_frames\": %d, \"frames_per_agent\": %d, \"rows\": [", track ? 1 : 0, pin ? 1 : 0, per);
    for (size_t ci = 0; ci < counts.size(); ++ci) {
        const int A = counts[ci];
        std::vector<Agent> agents(A);
        for (Agent& a : agents)
            if (!setup(a, device)) { fprintf(stderr, "%s\n", a.error.c_str()); return 5; }
        std::atomic<int> ready(0), bad(0);
        std::atomic<bool> go(false);
        std::vector<std::thread> th;
        for (int ai = 0; ai < A; ++ai)
            th.emplace_back([&, ai] {
                Agent& a = agents[ai];
                auto frame = [&](int i) { return frames.data() + (size_t)((ai * 3 + i) % nfile) * w * h; };
                for (int i = 0; i < 6; ++i)
                    if (!grab(a, frame(i), w, h, track)) { bad++; break; }
                ready++;
                while (!go.load(std::memory_order_acquire)) std::this_thread::yield();
                a.lat.reserve(per);
                a.tExtract = a.tUpdate = a.tSearch = 0;
                for (int i = 0; i < per && !bad.load(); ++i) {
                    Clock::time_point t0 = Clock::now();
                    if (!grab(a, frame(6 + i), w, h, track)) { bad++; break; }
                    a.lat.push_back(secondsSince(t0));
                }
            });
        while (ready.load() < A) std::this_thread::yield();
        Clock::time_point t0 = Clock::now();
        go.store(true, std::memory_order_release);
        for (std::thread& t : th) t.join();
        const double wall = secondsSince(t0);
        if (bad.load()) {
            for (Agent& a : agents) if (!a.error.empty()) fprintf(stderr, "%s\n", a.error.c_str());
            return 6;
        }
        std::vector<double> all;
        long long matches = 0, launches = 0;
        double tE = 0, tU = 0, tS = 0;
        for (Agent& a : agents) {
            tE += a.tExtract; tU += a.tUpdate; tS += a.tSearch;
            all.insert(all.end(), a.lat.begin(), a.lat.end());
            matches += a.matches;
            launches += orbx_launch_count(a.ex) + orbm_launch_count(a.mt);
        }
        printf("%s{\"agents\": %d, \"frames_per_s\": %.1f, \"median_frame_latency_ms\": %.4f, \"p95_frame_latency_ms\": %.4f, "
               "\"mean_call_ms\": {\"orbx_extract\": %.4f, \"orbm_frame_update\": %.4f, \"orbm_search_frame\": %.4f}, "
               "\"matches_per_frame\": %.1f, \"gpu_launches\": %lld}",
               ci ? ", " : "", A, (double)A * per / wall, percentile(all, 0.5) * 1e3, percentile(all, 0.95) * 1e3,
               tE * 1e3 / ((double)A * per), tU * 1e3 / ((double)A * per), tS * 1e3 / ((double)A * per), (double)matches / ((double)A * (per + 5)), launches);
        for (Agent& a : agents) {
            if (a.slot) orbm_frame_destroy(a.mt, a.slot);
            orbm_destroy(a.mt);
            orbx_destroy(a.ex);
        }
    }
    printf("]}\n");
    if (pin) orbx_unpin_host(frames.data());
    return 0;
}
