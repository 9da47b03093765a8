// Multi-replica MRMW register of fixed-size byte buffers with optional blocking
// producer/consumer ("consume") semantics, behind a C ABI.
//
// Role parity: /root/reference/pytorch_impl/libs/native/include/multiregister.hpp:232-552
// (ring of replicas in states empty/writing/written/readable/reading, blocking read()/write()
// with optional timeout, close()). In this framework it backs
// garfield_b200.utils.register.MultiRegister, used to hand the latest aggregated gradient /
// model between the step loop and serving threads without copying under a lock.
//
// Design: one mutex + two condition variables; each replica has a state, a reader count and a
// sequence number. Writers take any replica that is empty, or published-but-unread-by-anyone
// and older than the newest one; readers take the newest published replica (in consume mode a
// published replica is handed to exactly one reader, oldest first).

#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <vector>

#include "garfield_b200.h"

namespace {

enum class State : std::uint8_t { empty, writing, published, reading };

struct Replica {
    State state = State::empty;
    std::uint32_t readers = 0;
    std::uint64_t seq = 0;
    std::vector<unsigned char> bytes;
};

struct Register {
    std::mutex mutex;
    std::condition_variable can_read, can_write;
    std::vector<Replica> ring;
    std::uint64_t next_seq = 1;
    bool consume = false, closed = false;
    std::size_t size = 0;
};

bool wait_on(std::condition_variable& cv, std::unique_lock<std::mutex>& lock, std::int64_t timeout_ns) {
    if (timeout_ns < 0) {
        cv.wait(lock);
        return true;
    }
    return cv.wait_for(lock, std::chrono::nanoseconds(timeout_ns)) != std::cv_status::timeout;
}

int pick_write(Register& r) {
    int newest = -1, stale = -1;
    for (int i = 0; i < static_cast<int>(r.ring.size()); ++i) {
        Replica& rep = r.ring[i];
        if (rep.state == State::empty) return i;
        if (rep.state == State::published && (newest < 0 || rep.seq > r.ring[newest].seq)) newest = i;
    }
    if (r.consume) return -1;  // consume mode never drops unread data
    for (int i = 0; i < static_cast<int>(r.ring.size()); ++i) {
        Replica& rep = r.ring[i];
        if (rep.state == State::published && rep.readers == 0 && i != newest &&
            (stale < 0 || rep.seq < r.ring[stale].seq))
            stale = i;
    }
    if (stale < 0 && r.ring.size() == 1 && newest >= 0 && r.ring[newest].readers == 0) stale = newest;
    return stale;
}

int pick_read(Register& r) {
    int pick = -1;
    for (int i = 0; i < static_cast<int>(r.ring.size()); ++i) {
        Replica& rep = r.ring[i];
        if (rep.state != State::published) continue;
        if (pick < 0) pick = i;
        else if (r.consume ? rep.seq < r.ring[pick].seq : rep.seq > r.ring[pick].seq) pick = i;
    }
    return pick;
}

}  // namespace

extern "C" {

GB_API void* gb_mreg_create(int replicas, int64_t nbytes, int consume) {
    if (replicas < 1 || nbytes < 0) return nullptr;
    auto* r = new Register();
    r->ring.resize(static_cast<std::size_t>(replicas));
    for (auto& rep : r->ring) rep.bytes.resize(static_cast<std::size_t>(nbytes));
    r->consume = consume != 0;
    r->size = static_cast<std::size_t>(nbytes);
    return r;
}

GB_API void gb_mreg_close(void* handle) {
    auto* r = static_cast<Register*>(handle);
    {
        std::lock_guard<std::mutex> lock(r->mutex);
        r->closed = true;
    }
    r->can_read.notify_all();
    r->can_write.notify_all();
}

GB_API void gb_mreg_destroy(void* handle) {
    gb_mreg_close(handle);
    delete static_cast<Register*>(handle);
}

// Returns 0 ok, 1 closed, 2 timed out.
GB_API int gb_mreg_write(void* handle, void const* src, int64_t timeout_ns) {
    auto* r = static_cast<Register*>(handle);
    std::unique_lock<std::mutex> lock(r->mutex);
    int slot;
    for (;;) {
        if (r->closed) return 1;
        slot = pick_write(*r);
        if (slot >= 0) break;
        if (!wait_on(r->can_write, lock, timeout_ns)) return 2;
    }
    Replica& rep = r->ring[slot];
    rep.state = State::writing;
    lock.unlock();
    std::memcpy(rep.bytes.data(), src, r->size);  // copy outside the lock: the slot is ours
    lock.lock();
    rep.seq = r->next_seq++;
    rep.readers = 0;
    rep.state = State::published;
    lock.unlock();
    r->can_read.notify_all();
    return 0;
}

// Returns 0 ok (dst filled, *seq set), 1 closed, 2 timed out.
GB_API int gb_mreg_read(void* handle, void* dst, uint64_t* seq, int64_t timeout_ns) {
    auto* r = static_cast<Register*>(handle);
    std::unique_lock<std::mutex> lock(r->mutex);
    int slot;
    for (;;) {
        slot = pick_read(*r);
        if (slot >= 0) break;
        if (r->closed) return 1;
        if (!wait_on(r->can_read, lock, timeout_ns)) return 2;
    }
    Replica& rep = r->ring[slot];
    ++rep.readers;
    if (r->consume) rep.state = State::reading;  // no further reader may join
    lock.unlock();
    std::memcpy(dst, rep.bytes.data(), r->size);
    if (seq) *seq = rep.seq;
    lock.lock();
    --rep.readers;
    if (r->consume) {
        rep.state = State::empty;
        lock.unlock();
        r->can_write.notify_all();
    } else {
        bool const idle = rep.readers == 0;
        lock.unlock();
        if (idle) r->can_write.notify_all();
    }
    return 0;
}

}  // extern "C"
