// Test program for lidarslam_b200/csrc/host/worker.h (the helper threads behind the deferred map update and the staged
// inputs): jobs run in posting order, wait(ticket) returns once that job is over and rethrows what it threw, done() does
// not block, drain() swallows the errors nobody waited for, sleeping helpers wake up, destruction joins.  No GPU needed
// (cudaSetDevice may fail in the helper; nothing here touches the device).
#include "../../lidarslam_b200/csrc/host/worker.h"
#include <cstdio>
#include <stdexcept>
#include <vector>

#define CHECK(x) do { if (!(x)) { std::printf("FAILED %s:%d %s\n", __FILE__, __LINE__, #x); return 1; } } while (0)

int main()
{
  for (int spin = 0; spin < 2; ++spin) {
    lsb::AsyncWorker w(0);
    std::vector<int> order;
    std::atomic<int> slow{0};
    unsigned t[64];
    for (int i = 0; i < 64; ++i)
      t[i] = w.post([&order, &slow, i] {
        if (i == 3) { std::this_thread::sleep_for(std::chrono::milliseconds(30)); slow.store(1); }
        order.push_back(i);
      }, spin != 0);
    CHECK(!w.done(t[63]));                     // job 3 sleeps: the queue cannot be through yet
    w.wait(t[10]);
    CHECK(slow.load() == 1 && w.done(t[3]) && w.done(t[10]));
    w.wait(t[63]);
    CHECK(order.size() == 64);
    for (int i = 0; i < 64; ++i) CHECK(order[i] == i);
    // an exception travels to the thread that waits for that job, and to nobody else
    const unsigned bad = w.post([] { throw std::runtime_error("boom"); }, spin != 0);
    const unsigned good = w.post([] {}, spin != 0);
    w.wait(good);
    bool caught = false;
    try { w.wait(bad); } catch (const std::runtime_error& e) { caught = std::string(e.what()) == "boom"; }
    CHECK(caught);
    w.wait(bad);                               // reported once
    // an idle helper sleeps (at once with spin == 0, after its linger otherwise) and wakes up for the next job
    std::this_thread::sleep_for(std::chrono::milliseconds(spin ? 60 : 5));
    std::atomic<int> woke{0};
    w.wait(w.post([&woke] { woke.store(1); }, spin != 0));
    CHECK(woke.load() == 1);
    // drain: everything posted so far is over, errors nobody waited for are dropped
    for (int i = 0; i < 8; ++i) w.post([i] { if (i == 5) throw std::runtime_error("ignored"); }, spin != 0);
    w.drain();
    CHECK(w.done(w.last_posted()));
  }
  {  // destruction with jobs behind it: joins without hanging
    lsb::AsyncWorker w(0);
    for (int i = 0; i < 4; ++i) w.post([] { std::this_thread::sleep_for(std::chrono::milliseconds(2)); });
    w.drain();
  }
  std::printf("worker ok\n");
  return 0;
}
