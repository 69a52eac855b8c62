// TEST INFRASTRUCTURE ONLY (part of oracle/): a minimal stand-in for the slice of Intel TBB
// that the reference's PS_SimdPoly uses (reference: Parsip100/PS_SimdPoly/include/PS_Polygonizer.h:4-8,
// PS_Polygonizer.cpp:15-19,33-50,377-382,441-470). TBB is not installed in this image and only
// schedules work; no arithmetic lives in it. The shim is genuinely parallel: parallel_for spawns
// worker threads, each worker first calls every registered task_scheduler_observer (so the
// reference's ThreadStartSetup makes its per-thread replica), then pulls chunks off an atomic
// counter. Thread count: tbbshim::set_num_threads() or env PS_ORACLE_THREADS, default = all cores.
#ifndef TBBSHIM_CORE_H
#define TBBSHIM_CORE_H
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>   // real TBB headers pull these in transitively; the 2011-era reference relies on it
#include <cstdio>
#include <map>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>
#include <pthread.h>

namespace tbb {

class task_scheduler_observer;

namespace shim {
inline std::mutex& registry_mutex() { static std::mutex m; return m; }
inline std::vector<task_scheduler_observer*>& registry() { static std::vector<task_scheduler_observer*> r; return r; }
inline int& thread_override() { static int n = 0; return n; }
inline int num_threads() {
    if (thread_override() > 0) return thread_override();
    const char* e = std::getenv("PS_ORACLE_THREADS");
    if (e && std::atoi(e) > 0) return std::atoi(e);
    unsigned hc = std::thread::hardware_concurrency();
    return hc ? (int)hc : 1;
}
// A warm worker pool, like the arena real TBB keeps between parallel algorithms (the first version of this shim created and joined
// its pthreads in every parallel_for, which charged thread start-up - 256 MB stacks - to every Polygonize of the reference).
// run(n, fn, arg): n workers call fn(arg) concurrently, the caller waits. Workers are created on demand and never exit.
class Pool {
public:
    static Pool& get() { static Pool* p = new Pool(); return *p; }  // leaked on purpose: workers outlive static destruction
    static bool& in_worker() { static thread_local bool w = false; return w; }
    void run(int n, void (*fn)(void*), void* arg) {
        if (n <= 0) return;
        if (in_worker()) { fn(arg); return; }  // nested parallelism: serial (the reference does not nest)
        std::lock_guard<std::mutex> one(m_callers);  // one parallel algorithm at a time
        std::unique_lock<std::mutex> lk(m_m);
        while ((int)m_threads.size() < n) {
            pthread_attr_t attr; pthread_attr_init(&attr);
            pthread_attr_setstacksize(&attr, (size_t)256 << 20);  // widened reference builds keep MAX_TREE_NODES-sized scratch arrays on the stack
            Arg* a = new Arg{this, (int)m_threads.size()};
            pthread_t t;
            if (pthread_create(&t, &attr, &Pool::worker, a) != 0) { delete a; pthread_attr_destroy(&attr); break; }
            pthread_attr_destroy(&attr);
            m_threads.push_back(t);
        }
        if ((int)m_threads.size() < n) n = (int)m_threads.size();
        if (n == 0) { lk.unlock(); fn(arg); return; }
        m_fn = fn; m_arg = arg; m_want = n; m_running = n; m_epoch++;
        m_work.notify_all();
        m_done.wait(lk, [this] { return m_running == 0; });
    }
private:
    struct Arg { Pool* pool; int idx; };
    static void* worker(void* p) {
        Arg* a = (Arg*)p;
        Pool* P = a->pool;
        const int idx = a->idx;
        delete a;
        in_worker() = true;
        unsigned long long seen = 0;
        std::unique_lock<std::mutex> lk(P->m_m);
        for (;;) {
            P->m_work.wait(lk, [&] { return P->m_epoch != seen; });
            seen = P->m_epoch;
            if (idx >= P->m_want) continue;
            void (*fn)(void*) = P->m_fn; void* arg = P->m_arg;
            lk.unlock();
            fn(arg);
            lk.lock();
            if (--P->m_running == 0) P->m_done.notify_all();
        }
        return 0;
    }
    std::mutex m_m, m_callers;
    std::condition_variable m_work, m_done;
    std::vector<pthread_t> m_threads;
    void (*m_fn)(void*) = 0; void* m_arg = 0;
    int m_want = 0, m_running = 0;
    unsigned long long m_epoch = 0;
};
}  // namespace shim

class task_scheduler_observer {
public:
    task_scheduler_observer() : m_observing(false) {}
    virtual ~task_scheduler_observer() { observe(false); }
    void observe(bool state = true) {
        std::lock_guard<std::mutex> g(shim::registry_mutex());
        std::vector<task_scheduler_observer*>& r = shim::registry();
        if (state && !m_observing) { r.push_back(this); m_observing = true; }
        else if (!state && m_observing) {
            for (size_t i = 0; i < r.size(); i++) if (r[i] == this) { r.erase(r.begin() + i); break; }
            m_observing = false;
        }
    }
    virtual void on_scheduler_entry(bool) {}
    virtual void on_scheduler_exit(bool) {}
private:
    bool m_observing;
};

template <typename T>
class blocked_range {
public:
    typedef T const_iterator;
    blocked_range(T b, T e, size_t grain = 1) : m_b(b), m_e(e), m_grain(grain) {}
    T begin() const { return m_b; }
    T end() const { return m_e; }
    size_t size() const { return (size_t)(m_e - m_b); }
    size_t grainsize() const { return m_grain; }
    bool empty() const { return !(m_b < m_e); }
private:
    T m_b, m_e; size_t m_grain;
};

class auto_partitioner {};
class simple_partitioner {};

template <typename Range, typename Body>
void parallel_for(const Range& range, const Body& body) {
    const size_t n = range.size();
    if (n == 0) return;
    int nt = shim::num_threads();
    if ((size_t)nt > n) nt = (int)n;
    // chunk so that every worker gets ~16 chunks (work stealing stand-in)
    size_t chunk = n / ((size_t)nt * 16);
    if (chunk < 1) chunk = 1;
    std::atomic<size_t> next(0);
    std::vector<task_scheduler_observer*> obs;
    { std::lock_guard<std::mutex> g(shim::registry_mutex()); obs = shim::registry(); }
    struct Ctx { const Range* range; const Body* body; std::atomic<size_t>* next; size_t chunk, n;
                 std::vector<task_scheduler_observer*>* obs; };
    Ctx ctx = { &range, &body, &next, chunk, n, &obs };
    struct Run {
        static void go(void* p) {
            Ctx* c = (Ctx*)p;
            for (size_t i = 0; i < c->obs->size(); i++) (*c->obs)[i]->on_scheduler_entry(true);
            for (;;) {
                size_t b = c->next->fetch_add(c->chunk);
                if (b >= c->n) break;
                size_t e = b + c->chunk; if (e > c->n) e = c->n;
                Range sub(c->range->begin() + b, c->range->begin() + e);
                (*c->body)(sub);
            }
        }
    };
    shim::Pool::get().run(nt, &Run::go, &ctx);
}
template <typename Range, typename Body, typename Part>
void parallel_for(const Range& range, const Body& body, const Part&) { parallel_for(range, body); }

template <typename T>
class enumerable_thread_specific {
    typedef std::map<std::thread::id, std::unique_ptr<T> > map_t;
public:
    typedef T& reference;
    class const_iterator {
    public:
        const_iterator(typename map_t::const_iterator it) : m_it(it) {}
        const T& operator*() const { return *m_it->second; }
        const T* operator->() const { return m_it->second.get(); }
        const_iterator& operator++() { ++m_it; return *this; }
        const_iterator operator++(int) { const_iterator t(*this); ++m_it; return t; }
        bool operator!=(const const_iterator& o) const { return m_it != o.m_it; }
        bool operator==(const const_iterator& o) const { return m_it == o.m_it; }
    private:
        typename map_t::const_iterator m_it;
    };
    enumerable_thread_specific() : m_hasExemplar(false) {}
    explicit enumerable_thread_specific(const T& ex) : m_hasExemplar(true), m_exemplar(new T(ex)) {}
    reference local() {
        std::lock_guard<std::mutex> g(m_mutex);
        std::unique_ptr<T>& p = m_map[std::this_thread::get_id()];
        if (!p) p.reset(m_hasExemplar ? new T(*m_exemplar) : new T());
        return *p;
    }
    void clear() { std::lock_guard<std::mutex> g(m_mutex); m_map.clear(); }
    const_iterator begin() const { return const_iterator(m_map.begin()); }
    const_iterator end() const { return const_iterator(m_map.end()); }
    size_t size() const { return m_map.size(); }
private:
    std::mutex m_mutex;
    map_t m_map;
    bool m_hasExemplar;
    std::unique_ptr<T> m_exemplar;
};

class tick_count {
public:
    class interval_t {
    public:
        explicit interval_t(double s = 0) : m_s(s) {}
        double seconds() const { return m_s; }
    private:
        double m_s;
    };
    tick_count() {}
    static tick_count now() { tick_count t; t.m_t = std::chrono::steady_clock::now(); return t; }
    friend interval_t operator-(const tick_count& a, const tick_count& b) {
        return interval_t(std::chrono::duration<double>(a.m_t - b.m_t).count());
    }
private:
    std::chrono::steady_clock::time_point m_t;
};

struct tbb_thread { typedef std::thread::id id; };
namespace this_tbb_thread { inline std::thread::id get_id() { return std::this_thread::get_id(); } }

}  // namespace tbb

namespace tbbshim { inline void set_num_threads(int n) { tbb::shim::thread_override() = n; } }
#endif
