// ThreadSanitizer harness for the input-pipeline runtime (csrc/host/loader.cpp): worker threads fill a ring of slots
// while the consumer takes, checks and releases batches; whole epochs (every sample exactly once, every batch in its
// slot, a new order each epoch), epochs abandoned half way (steps_per_epoch / evaluate(steps=...)), a consumer that
// holds several slots before releasing them, and destruction while workers wait. Built by tests/test_cpu_loader.py
// with -fsanitize=thread: a data race in the slot protocol is a TSan report and a non-zero exit.
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <chrono>
#include <vector>

#include "idcb200_host.h"

#define CHECK(c) do { if (!(c)) { std::printf("CHECK failed: %s (line %d)\n", #c, __LINE__); return 1; } } while (0)

int main() {
    const long long n = 1003, sf = 24;
    const int batch = 32, slots = 4, threads = 4;
    std::vector<float> images(static_cast<size_t>(n * sf)), labels(static_cast<size_t>(n));
    for (long long i = 0; i < n; ++i) {
        labels[static_cast<size_t>(i)] = static_cast<float>(i);
        for (long long j = 0; j < sf; ++j) images[static_cast<size_t>(i * sf + j)] = static_cast<float>(i * 100 + j);
    }
    std::vector<std::vector<float>> sx(slots, std::vector<float>(static_cast<size_t>(batch * sf))),
        sy(slots, std::vector<float>(static_cast<size_t>(batch)));
    std::vector<float*> px, py;
    for (int s = 0; s < slots; ++s) { px.push_back(sx[static_cast<size_t>(s)].data()); py.push_back(sy[static_cast<size_t>(s)].data()); }
    void* h = idcbh_loader_create(images.data(), labels.data(), n, sf, batch, 100, 0, slots, px.data(), py.data(), threads, 5);
    CHECK(h != nullptr);
    CHECK(idcbh_loader_num_batches(h) == 32 || idcbh_loader_num_batches(h) == 0);   // (defined after the first start_epoch)
    std::vector<float> first_epoch_order;
    for (int epoch = 0; epoch < 6; ++epoch) {
        CHECK(idcbh_loader_start_epoch(h) == 0);
        CHECK(idcbh_loader_num_batches(h) == 32);
        std::vector<int> seen(static_cast<size_t>(n), 0);
        std::vector<float> order;
        std::vector<int> held;
        const bool abandon = epoch == 2 || epoch == 4;       // stop early, like evaluate(steps=...)
        int slot = -1, count = 0;
        long long k = 0;
        while (true) {
            const int r = idcbh_loader_next(h, &slot, &count);
            CHECK(r >= 0);
            if (r == 0) break;
            CHECK(slot == static_cast<int>(k % slots));
            CHECK(count == (k == 31 ? 1003 - 31 * 32 : 32));
            for (int i = 0; i < count; ++i) {
                const float id = sy[static_cast<size_t>(slot)][static_cast<size_t>(i)];
                const long long s = static_cast<long long>(id);
                CHECK(s >= 0 && s < n);
                seen[static_cast<size_t>(s)]++;
                order.push_back(id);
                for (long long j = 0; j < sf; j += 7)
                    CHECK(sx[static_cast<size_t>(slot)][static_cast<size_t>(i * sf + j)] == static_cast<float>(s * 100 + j));
            }
            ++k;
            // hold up to two slots before giving them back (a consumer with a copy in flight), sometimes slowly
            held.push_back(slot);
            if (held.size() >= 2) { CHECK(idcbh_loader_release(h, held.front()) == 0); held.erase(held.begin()); }
            if (k % 5 == 0) std::this_thread::sleep_for(std::chrono::microseconds(200));
            if (abandon && k == 11) break;
        }
        for (int s : held) CHECK(idcbh_loader_release(h, s) == 0);
        if (!abandon) {
            for (long long i = 0; i < n; ++i) CHECK(seen[static_cast<size_t>(i)] == 1);
            if (first_epoch_order.empty()) first_epoch_order = order;
            else CHECK(order != first_epoch_order);          // reshuffled each epoch
        }
    }
    idcbh_loader_destroy(h);
    // destruction right after creation and in the middle of an epoch (workers waiting / filling)
    h = idcbh_loader_create(images.data(), labels.data(), n, sf, batch, 100, 1, slots, px.data(), py.data(), threads, 6);
    CHECK(h != nullptr);
    idcbh_loader_destroy(h);
    h = idcbh_loader_create(images.data(), labels.data(), n, sf, batch, 1, 1, slots, px.data(), py.data(), threads, 7);
    CHECK(h != nullptr && idcbh_loader_start_epoch(h) == 0);
    CHECK(idcbh_loader_num_batches(h) == 31);               // drop_remainder
    int slot = 0, count = 0;
    CHECK(idcbh_loader_next(h, &slot, &count) == 1 && count == 32);
    CHECK(sy[static_cast<size_t>(slot)][0] == 0.f && sy[static_cast<size_t>(slot)][31] == 31.f);   // shuffle buffer 1 = input order
    idcbh_loader_destroy(h);
    std::printf("loader ok\n");
    return 0;
}
