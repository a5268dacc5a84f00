// ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// std::-based stand-ins for the handful of libcore runtime symbols the reference kd-tree
// builder links against but which cannot be compiled here (they need boost::thread,
// boost::filesystem, ...): Mutex, ConditionVariable, Thread, Logger::log, allocAligned,
// formatString, memString, getCoreCount.  Declarations come from the reference headers;
// the bodies are new and only serve oracle/_ref/libmtsref.so.
#include <mitsuba/mitsuba.h>
#include <mitsuba/core/lock.h>
#include <mitsuba/core/thread.h>
#include <mitsuba/core/logger.h>
#include <mitsuba/core/random.h>
#include <mutex>
#include <condition_variable>
#include <thread>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>

MTS_NAMESPACE_BEGIN

// ---- memory / strings ----------------------------------------------------
void * __restrict allocAligned(size_t size) {
    void *p = NULL;
    if (posix_memalign(&p, 64, size ? size : 1) != 0) throw std::bad_alloc();
    return p;
}
void freeAligned(void *ptr) { free(ptr); }

std::string formatString(const char *fmt, ...) {
    char tmp[2048];
    va_list ap; va_start(ap, fmt);
    int n = vsnprintf(tmp, sizeof(tmp), fmt, ap);
    va_end(ap);
    if (n < (int) sizeof(tmp)) return std::string(tmp);
    std::string big((size_t) n + 1, '\0');
    va_start(ap, fmt); vsnprintf(&big[0], big.size(), fmt, ap); va_end(ap);
    big.resize((size_t) n);
    return big;
}
std::string memString(size_t size, bool) {
    const char *unit[] = {"B", "KiB", "MiB", "GiB", "TiB"};
    double v = (double) size; int i = 0;
    while (v >= 1024 && i < 4) { v /= 1024; ++i; }
    return formatString("%.2f %s", v, unit[i]);
}
static int g_cores = 0;
int getCoreCount() { return g_cores > 0 ? g_cores : (int) std::max(1u, std::thread::hardware_concurrency()); }

// ---- Mutex / ConditionVariable ---------------------------------------------
struct Mutex::MutexPrivate { std::recursive_mutex m; };
Mutex::Mutex() : d(new MutexPrivate()) {}
Mutex::~Mutex() {}
void Mutex::lock() { d->m.lock(); }
void Mutex::unlock() { d->m.unlock(); }

struct ConditionVariable::ConditionVariablePrivate {
    ref<Mutex> mutex;
    std::condition_variable_any cv;
};
ConditionVariable::ConditionVariable(Mutex *mutex) : d(new ConditionVariablePrivate()) {
    d->mutex = mutex != NULL ? mutex : new Mutex();
}
ConditionVariable::~ConditionVariable() {}
void ConditionVariable::signal() { d->cv.notify_one(); }
void ConditionVariable::broadcast() { d->cv.notify_all(); }
void ConditionVariable::wait() { d->cv.wait(d->mutex->d->m); }
bool ConditionVariable::wait(int ms) {
    if (ms == -1) { wait(); return true; }
    return d->cv.wait_for(d->mutex->d->m, std::chrono::milliseconds(ms)) == std::cv_status::no_timeout;
}

// ---- Logger (only log() is reachable) ---------------------------------------
static int g_logLevel = EWarn;
void Logger::log(ELogLevel level, const Class *theClass, const char *file, int line, const char *fmt, ...) {
    if (level < g_logLevel && level != EError) return;
    char tmp[4096];
    va_list ap; va_start(ap, fmt); vsnprintf(tmp, sizeof(tmp), fmt, ap); va_end(ap);
    if (level == EError) throw std::runtime_error(std::string("[reference] ") + tmp);
    fprintf(stderr, "[reference%s%s] %s\n", theClass ? " " : "", theClass ? theClass->getName().c_str() : "", tmp);
}

// ---- Thread --------------------------------------------------------------------
struct Thread::ThreadPrivate {
    std::string name;
    std::thread th;
    bool running;
    ThreadPrivate(const std::string &n) : name(n), running(false) {}
};
static thread_local Thread *t_self = NULL;
alignas(64) static char g_fakeLogger[256];      // Logger::log above never touches `this`

namespace {
class MainThread : public Thread {
public:
    MainThread() : Thread("main") {}
    void run() {}
};
}

Thread::Thread(const std::string &name) : d(new ThreadPrivate(name)) {}
Thread::~Thread() { if (d->th.joinable()) d->th.join(); }
void Thread::setCritical(bool) {}
Logger *Thread::getLogger() { return reinterpret_cast<Logger *>(g_fakeLogger); }
Thread *Thread::getThread() {
    if (t_self == NULL) { t_self = new MainThread(); t_self->incRef(); }
    return t_self;
}
void Thread::dispatch(Thread *thread) { t_self = thread; thread->run(); thread->d->running = false; }
void Thread::start() {
    d->running = true;
    incRef();
    d->th = std::thread([this]() { Thread::dispatch(this); });
}
void Thread::join() { if (d->th.joinable()) d->th.join(); decRef(false); }
std::string Thread::toString() const { return "Thread[" + d->name + "]"; }

// util.cpp's stratified-sampling helpers take a Random; nothing on the paths used here calls them
Float Random::nextFloat() { throw std::runtime_error("[reference] Random is not part of this build"); }
size_t Random::nextSize(size_t) { throw std::runtime_error("[reference] Random is not part of this build"); }

MTS_IMPLEMENT_CLASS(Mutex, false, Object)
MTS_IMPLEMENT_CLASS(ConditionVariable, false, Object)
MTS_IMPLEMENT_CLASS(Thread, true, Object)

MTS_NAMESPACE_END

extern "C" {
void ref_static_init() {
    static std::once_flag once;
    std::call_once(once, []() { mitsuba::Class::staticInitialization(); });
}
void ref_set_log_level(int level) { mitsuba::g_logLevel = level; }
void ref_set_core_count(int n) { mitsuba::g_cores = n; }
}
