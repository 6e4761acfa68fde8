// TEST INFRASTRUCTURE ONLY -- the part of roscpp the reference nodes touch (ekf.cpp:38-46,74-99,313-326 and the
// same lines of ukf.cpp / sensor_landmark.cpp), as an in-process stand-in: subscribe() records the member callback,
// advertise().publish() keeps the last message, ros::Time::now() is a settable clock. One Context per node object,
// selected through a thread-local pointer so that many reference nodes can run on OpenMP threads (the CPU baseline).
#pragma once
#include <cstdint>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace ros {
namespace shim {

struct Context {
    double now = 0.0;
    // topic -> callback taking a pointer to a std::shared_ptr<const Msg>
    std::map<std::string, std::function<void(const void*)>> subscribers;
    // topic -> last published message (type-erased copy) and publish count
    std::map<std::string, std::shared_ptr<void>> last;
    std::map<std::string, uint64_t> published;

    template <class M>
    bool dispatch(const std::string& topic, const std::shared_ptr<const M>& msg) {
        auto it = subscribers.find(topic);
        if (it == subscribers.end()) return false;
        it->second(&msg);
        return true;
    }
    template <class M>
    const M* last_published(const std::string& topic) const {
        auto it = last.find(topic);
        return it == last.end() ? nullptr : static_cast<const M*>(it->second.get());
    }
};

inline Context*& current() {
    static thread_local Context* c = nullptr;
    return c;
}

// RAII selection of the context a node is constructed in / dispatched under.
struct Use {
    Context* prev;
    explicit Use(Context* c) : prev(current()) { current() = c; }
    ~Use() { current() = prev; }
};

}  // namespace shim

struct Time {
    double t = 0.0;
    static void init() {}
    static Time now() {
        Time x;
        x.t = shim::current() ? shim::current()->now : 0.0;
        return x;
    }
    double toSec() const { return t; }
};

struct Rate {
    explicit Rate(double) {}
    void sleep() {}
};

struct Subscriber {};

struct Publisher {
    shim::Context* ctx = nullptr;
    std::string topic;
    template <class M>
    void publish(const M& msg) const {
        if (!ctx) return;
        ctx->last[topic] = std::make_shared<M>(msg);
        ctx->published[topic] += 1;
    }
};

struct NodeHandle {
    shim::Context* ctx;
    NodeHandle() : ctx(shim::current()) {}
    template <class M, class T>
    Subscriber subscribe(const std::string& topic, uint32_t, void (T::*fp)(const std::shared_ptr<const M>&), T* obj) {
        if (ctx) ctx->subscribers[topic] = [fp, obj](const void* p) { (obj->*fp)(*static_cast<const std::shared_ptr<const M>*>(p)); };
        return Subscriber();
    }
    template <class M>
    Publisher advertise(const std::string& topic, uint32_t) {
        Publisher p;
        p.ctx = ctx;
        p.topic = topic;
        return p;
    }
};

inline void init(int&, char**, const std::string&) {}
inline bool ok() { return false; }  // the reference's main() loops while ros::ok(): it returns at once here
inline void spinOnce() {}

}  // namespace ros
