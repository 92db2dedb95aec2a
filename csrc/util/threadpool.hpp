// Fixed-size worker pool with a bounded FIFO.  Replaces the reference's pi::threadpool
// (reference: third_party/threadpool/include/pithreadpool/threadpool.hpp:192-299): tasks are
// std::function<void()>, completion is reported through std::future.
#pragma once
#include <condition_variable>
#include <deque>
#include <functional>
#include <future>
#include <mutex>
#include <thread>
#include <vector>

namespace pccl {

class ThreadPool {
public:
    ThreadPool(size_t n_threads, size_t max_queue) : max_queue_(max_queue) {
        for (size_t i = 0; i < n_threads; ++i) workers_.emplace_back([this] { loop(); });
    }
    ~ThreadPool() { shutdown(); }
    ThreadPool(const ThreadPool &) = delete;
    ThreadPool &operator=(const ThreadPool &) = delete;

    // Blocks while the queue is full.
    template <typename F> auto submit(F &&fn) -> std::future<decltype(fn())> {
        using R = decltype(fn());
        auto task = std::make_shared<std::packaged_task<R()>>(std::forward<F>(fn));
        auto fut = task->get_future();
        {
            std::unique_lock<std::mutex> lock(mu_);
            cv_space_.wait(lock, [this] { return stop_ || queue_.size() < max_queue_; });
            if (stop_) return fut; // broken promise -> future throws; callers treat as failure
            queue_.emplace_back([task] { (*task)(); });
        }
        cv_work_.notify_one();
        return fut;
    }

    void shutdown() {
        {
            std::lock_guard<std::mutex> lock(mu_);
            if (stop_) return;
            stop_ = true;
        }
        cv_work_.notify_all();
        cv_space_.notify_all();
        for (auto &t : workers_)
            if (t.joinable()) t.join();
    }

private:
    void loop() {
        for (;;) {
            std::function<void()> job;
            {
                std::unique_lock<std::mutex> lock(mu_);
                cv_work_.wait(lock, [this] { return stop_ || !queue_.empty(); });
                if (queue_.empty()) return;
                job = std::move(queue_.front());
                queue_.pop_front();
            }
            cv_space_.notify_one();
            job();
        }
    }

    size_t max_queue_;
    std::mutex mu_;
    std::condition_variable cv_work_, cv_space_;
    std::deque<std::function<void()>> queue_;
    std::vector<std::thread> workers_;
    bool stop_ = false;
};

} // namespace pccl
