// Input pipeline runtime: the hot loop of `cache() -> shuffle(buffer) -> batch() -> prefetch()` as the
// reference builds it with tf.data (prepare_for_training, /root/reference/dist_model_tf_mobile.py:47-65;
// SURVEY.md 2.A A4, 2.B B14, 2.C "tf.data C++ runtime"). The cached dataset is one contiguous float32 sample
// array + one label array owned by the caller; worker threads assemble shuffled batches into a ring of
// caller-provided slots (pinned host memory, so the consumer can start an asynchronous H2D copy straight out
// of a slot) while the consumer trains on earlier batches.
//
// Shuffle semantics are tf.data's: a buffer of `shuffle_buffer` elements is filled from the input order, each
// output element is drawn uniformly from the buffer and replaced by the next input element; a new order is
// drawn every epoch (reshuffle_each_iteration). The last batch may be short unless drop_remainder is set.
//
// Ordering protocol: batch k always goes to slot k % n_slots. Workers claim batch numbers in increasing order
// and wait for their slot to be FREE; the consumer takes batches in increasing order and waits for its slot
// to be READY. A slot becomes FREE again only when the consumer releases it, so the ring cannot overrun and
// cannot deadlock (the lowest unfinished batch always owns a slot that is either FREE or being filled).
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <random>
#include <thread>
#include <vector>

#include "host_common.h"
#include "idcb200_host.h"

namespace idcbh {

struct Loader {
    const float* images;
    const float* labels;
    long long n, sample_floats;
    int batch, shuffle_buffer, drop_remainder, n_slots;
    std::vector<float*> slot_x, slot_y;
    unsigned long long seed;
    unsigned long long epoch = 0;

    enum State : int { FREE = 0, FILLING = 1, READY = 2 };
    std::vector<int> state;
    std::vector<long long> slot_batch;
    std::vector<long long> order;      // this epoch's sample order
    long long n_batches = 0;
    long long next_fill = 0, next_consume = 0;
    bool stop = false;
    std::mutex mu;
    std::condition_variable cv_work, cv_ready;
    std::vector<std::thread> workers;

    void make_order() {
        // tf.data shuffle: reservoir of `shuffle_buffer` indices
        order.resize(static_cast<size_t>(n));
        const long long B = shuffle_buffer <= 1 ? 1 : (shuffle_buffer < n ? shuffle_buffer : n);
        std::mt19937_64 rng(seed * 0x9E3779B97F4A7C15ull + epoch + 1);
        if (B <= 1) {
            for (long long i = 0; i < n; ++i) order[static_cast<size_t>(i)] = i;
            return;
        }
        std::vector<long long> buf(static_cast<size_t>(B));
        for (long long i = 0; i < B; ++i) buf[static_cast<size_t>(i)] = i;
        long long next_in = B, filled = B, out = 0;
        while (filled > 0) {
            const long long j = static_cast<long long>(rng() % static_cast<unsigned long long>(filled));
            order[static_cast<size_t>(out++)] = buf[static_cast<size_t>(j)];
            if (next_in < n) {
                buf[static_cast<size_t>(j)] = next_in++;
            } else {
                buf[static_cast<size_t>(j)] = buf[static_cast<size_t>(--filled)];
            }
        }
    }

    long long batch_count(long long k) const {
        const long long lo = k * batch;
        const long long hi = lo + batch < n ? lo + batch : n;
        return hi - lo;
    }

    void work() {
        for (;;) {
            long long k;
            int s;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv_work.wait(lk, [&] {
                    return stop || (next_fill < n_batches && state[static_cast<size_t>(next_fill % n_slots)] == FREE);
                });
                if (stop) return;
                k = next_fill++;
                s = static_cast<int>(k % n_slots);
                state[static_cast<size_t>(s)] = FILLING;
                slot_batch[static_cast<size_t>(s)] = k;
            }
            const long long lo = k * batch, cnt = batch_count(k);
            float* dx = slot_x[static_cast<size_t>(s)];
            float* dy = slot_y[static_cast<size_t>(s)];
            const size_t bytes = static_cast<size_t>(sample_floats) * sizeof(float);
            for (long long i = 0; i < cnt; ++i) {
                const long long src = order[static_cast<size_t>(lo + i)];
                std::memcpy(dx + i * sample_floats, images + src * sample_floats, bytes);
                dy[i] = labels ? labels[src] : 0.f;
            }
            {
                std::lock_guard<std::mutex> lk(mu);
                state[static_cast<size_t>(s)] = READY;
            }
            cv_ready.notify_all();
        }
    }
};

}  // namespace idcbh

using idcbh::Loader;

extern "C" {

void* idcbh_loader_create(const float* images, const float* labels, long long n, long long sample_floats, int batch,
                          int shuffle_buffer, int drop_remainder, int n_slots, float* const* slot_x,
                          float* const* slot_y, int n_threads, unsigned long long seed) {
    if (!images || n < 0 || sample_floats <= 0 || batch <= 0 || n_slots < 2 || !slot_x || !slot_y || n_threads < 1) {
        idcbh::set_error("loader_create: bad arguments (need images, batch > 0, >= 2 slots, >= 1 thread)");
        return nullptr;
    }
    Loader* L = new Loader();
    L->images = images; L->labels = labels; L->n = n; L->sample_floats = sample_floats;
    L->batch = batch; L->shuffle_buffer = shuffle_buffer; L->drop_remainder = drop_remainder; L->n_slots = n_slots;
    L->seed = seed;
    for (int i = 0; i < n_slots; ++i) {
        if (!slot_x[i] || !slot_y[i]) {
            idcbh::set_error("loader_create: null slot buffer");
            delete L;
            return nullptr;
        }
        L->slot_x.push_back(slot_x[i]);
        L->slot_y.push_back(slot_y[i]);
    }
    L->state.assign(static_cast<size_t>(n_slots), Loader::FREE);
    L->slot_batch.assign(static_cast<size_t>(n_slots), -1);
    for (int t = 0; t < n_threads; ++t) L->workers.emplace_back([L] { L->work(); });
    return L;
}

// Begin a new epoch: draws a new order. Must not be called while slots of the previous epoch are still held.
int idcbh_loader_start_epoch(void* h) {
    Loader* L = static_cast<Loader*>(h);
    if (!L) return 1;
    {
        std::lock_guard<std::mutex> lk(L->mu);
        // batches of an abandoned epoch that are READY/FILLING must drain first: wait outside the lock below
    }
    {
        std::unique_lock<std::mutex> lk(L->mu);
        // stop further claims of the old epoch, then wait for in-flight fills
        L->n_batches = L->next_fill;
        L->cv_ready.wait(lk, [&] {
            for (int s : L->state)
                if (s == Loader::FILLING) return false;
            return true;
        });
        for (auto& s : L->state) s = Loader::FREE;
        L->make_order();
        L->epoch += 1;
        const long long full = L->n / L->batch, rem = L->n % L->batch;
        L->n_batches = full + ((rem && !L->drop_remainder) ? 1 : 0);
        L->next_fill = 0;
        L->next_consume = 0;
    }
    L->cv_work.notify_all();
    return 0;
}

// Blocks until the next batch of the epoch is ready. Returns 1 and (*slot, *count), or 0 at the end of the epoch.
int idcbh_loader_next(void* h, int* slot, int* count) {
    Loader* L = static_cast<Loader*>(h);
    if (!L || !slot || !count) return -1;
    std::unique_lock<std::mutex> lk(L->mu);
    const long long k = L->next_consume;
    if (k >= L->n_batches) return 0;
    const int s = static_cast<int>(k % L->n_slots);
    L->cv_ready.wait(lk, [&] { return L->state[static_cast<size_t>(s)] == Loader::READY && L->slot_batch[static_cast<size_t>(s)] == k; });
    L->next_consume = k + 1;
    *slot = s;
    *count = static_cast<int>(L->batch_count(k));
    return 1;
}

int idcbh_loader_release(void* h, int slot) {
    Loader* L = static_cast<Loader*>(h);
    if (!L || slot < 0 || slot >= L->n_slots) return 1;
    {
        std::lock_guard<std::mutex> lk(L->mu);
        L->state[static_cast<size_t>(slot)] = Loader::FREE;
    }
    L->cv_work.notify_all();
    return 0;
}

long long idcbh_loader_num_batches(void* h) {
    Loader* L = static_cast<Loader*>(h);
    return L ? L->n_batches : -1;
}

void idcbh_loader_destroy(void* h) {
    Loader* L = static_cast<Loader*>(h);
    if (!L) return;
    {
        std::lock_guard<std::mutex> lk(L->mu);
        L->stop = true;
    }
    L->cv_work.notify_all();
    for (auto& t : L->workers) t.join();
    delete L;
}

}  // extern "C"
