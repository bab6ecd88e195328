// CPU test of the stand-alone ringbuffer<T> used by the adapter, restating the reference's
// unittest/ringbuffer_test.cpp:13-177 cases: block size, 2-thread producer/consumer with 2 and 128
// slots, Stop() releases a blocked reader, peek*(-1) addresses the previous slot.
#include <cstdio>
#include <cstring>
#include <thread>

#include "r2iq_iface.h"

static int failures = 0;
#define CHECK(c) do { if (!(c)) { printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #c); failures++; } } while (0)

static void produce_consume(int slots) {
    ringbuffer<int16_t> rb(slots);
    rb.setBlockSize(1024);
    rb.Start();
    const int n = 1000;
    std::thread prod([&] {
        for (int i = 0; i < n; i++) {
            int16_t* p = rb.getWritePtr();
            for (int k = 0; k < 1024; k++) p[k] = (int16_t)(i + k);
            rb.WriteDone();
        }
    });
    long bad = 0;
    for (int i = 0; i < n; i++) {
        const int16_t* p = rb.getReadPtr();
        for (int k = 0; k < 1024; k++) bad += p[k] != (int16_t)(i + k);
        rb.ReadDone();
    }
    prod.join();
    CHECK(bad == 0);
    CHECK(rb.getWriteCount() == n);
}

int main() {
    {
        ringbuffer<int16_t> rb;
        rb.setBlockSize(128 * 1024);
        CHECK(rb.getBlockSize() == 128 * 1024);
        rb.setBlockSize(1021);              // rounded up internally, reported as asked
        CHECK(rb.getBlockSize() == 1021);
    }
    produce_consume(2);
    produce_consume(128);
    {   // Stop() releases a reader blocked on an empty ring
        ringbuffer<float> rb(4);
        rb.setBlockSize(64);
        rb.Start();
        bool returned = false;
        std::thread t([&] { (void)rb.getReadPtr(); returned = true; });
        std::this_thread::sleep_for(std::chrono::milliseconds(50));
        CHECK(!returned);
        rb.Stop();
        t.join();
        CHECK(returned);
    }
    {   // peekReadPtr(-1) is the slot consumed just before the current one
        ringbuffer<int16_t> rb(8);
        rb.setBlockSize(16);
        rb.Start();
        for (int i = 0; i < 3; i++) { int16_t* p = rb.getWritePtr(); p[0] = (int16_t)(100 + i); rb.WriteDone(); }
        (void)rb.getReadPtr(); rb.ReadDone();
        const int16_t* cur = rb.getReadPtr();
        CHECK(cur[0] == 101);
        CHECK(rb.peekReadPtr(-1)[0] == 100);
        CHECK(rb.peekWritePtr(-1)[0] == 102);
    }
    printf(failures ? "FAILED %d\n" : "ALL PASS\n", failures);
    return failures ? 1 : 0;
}
