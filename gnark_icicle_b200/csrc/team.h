// Host-side threading helpers of libg16b200: a team of persistent threads (multi.cu: one per GPU rank; HostStager: the
// copier threads that stage pageable caller memory) and the rendezvous the ranks use between their exchanges.
#pragma once
#include <condition_variable>
#include <cstdint>
#include <functional>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

namespace g16 {

// rendezvous of the rank threads; a failing rank releases everybody
struct HostBarrier {
    std::mutex m;
    std::condition_variable cv;
    int n = 1, waiting = 0;
    uint64_t gen = 0;
    bool failed = false;
    void reset(int n_) { std::lock_guard<std::mutex> lk(m); n = n_; waiting = 0; failed = false; }
    bool wait() {
        std::unique_lock<std::mutex> lk(m);
        if (failed) return false;
        const uint64_t g = gen;
        if (++waiting == n) { waiting = 0; gen++; cv.notify_all(); }
        else cv.wait(lk, [&] { return gen != g || failed; });
        return !failed;
    }
    void fail() { std::lock_guard<std::mutex> lk(m); failed = true; cv.notify_all(); }
};

// n - 1 persistent host threads; member 0 of every run() is the calling thread
struct Team {
    int n = 0;
    std::vector<std::thread> th;
    std::mutex m;
    std::condition_variable cv_go, cv_done;
    const std::function<void(int)>* job = nullptr;
    uint64_t gen = 0;
    int pending = 0;
    bool stop = false;
    std::vector<std::string> err;

    void start(int n_) {
        n = n_;
        err.assign(n, "");
        for (int r = 1; r < n; r++) th.emplace_back([this, r] { loop(r); });
    }
    void loop(int r) {
        uint64_t seen = 0;
        for (;;) {
            const std::function<void(int)>* f;
            {
                std::unique_lock<std::mutex> lk(m);
                cv_go.wait(lk, [&] { return stop || gen != seen; });
                if (stop) return;
                seen = gen;
                f = job;
            }
            try { (*f)(r); } catch (const std::exception& e) { err[r] = e.what(); if (err[r].empty()) err[r] = "error"; }
            {
                std::lock_guard<std::mutex> lk(m);
                pending--;
            }
            cv_done.notify_all();
        }
    }
    // f(r) for every rank, concurrently; throws the first rank's error
    void run(const std::function<void(int)>& f) {
        for (auto& e : err) e.clear();
        {
            std::lock_guard<std::mutex> lk(m);
            job = &f; gen++; pending = n - 1;
        }
        cv_go.notify_all();
        try { f(0); } catch (const std::exception& e) { err[0] = e.what(); if (err[0].empty()) err[0] = "error"; }
        {
            std::unique_lock<std::mutex> lk(m);
            cv_done.wait(lk, [&] { return pending == 0; });
        }
        for (int r = 0; r < n; r++)
            if (!err[r].empty() && err[r] != "another rank failed") throw std::runtime_error("rank " + std::to_string(r) + ": " + err[r]);
        for (int r = 0; r < n; r++)
            if (!err[r].empty()) throw std::runtime_error("rank " + std::to_string(r) + ": " + err[r]);
    }
    void shutdown() {
        {
            std::lock_guard<std::mutex> lk(m);
            stop = true;
        }
        cv_go.notify_all();
        for (auto& t : th) t.join();
        th.clear();
    }
};

}  // namespace g16
