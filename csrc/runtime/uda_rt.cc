#include "uda_rt.h"

#include <execinfo.h>
#include <pwd.h>
#include <unistd.h>

#include <atomic>
#include <cstdio>
#include <cstring>
#include <ctime>
#include <mutex>
#include <thread>

namespace uda {

// ------------------------------------------------------------------------------------------------ errors
std::string capture_backtrace(int skip) {
    void* frames[32];
    int n = ::backtrace(frames, 32);
    char** syms = ::backtrace_symbols(frames, n);
    std::string out;
    if (syms) {
        for (int i = skip; i < n; ++i) {
            out += "\n    at ";
            out += syms[i];
        }
        free(syms);
    }
    return out;
}

UdaError::UdaError(const std::string& info) : std::runtime_error(info), info_(info) {
    full_ = "UdaException: " + info + capture_backtrace(2);
}

// ------------------------------------------------------------------------------------------------ logging
namespace {
std::atomic<int> g_threshold{lsINFO};
std::atomic<bool> g_unique_file{false};
const HostCallbacks* g_host = nullptr;
FILE* g_file = nullptr;
std::mutex g_log_mu;
const char* kSeverityName[] = {"NONE", "FATAL", "ERROR", "WARN", "INFO", "DEBUG", "TRACE"};

void open_file(const std::string& path) {
    std::lock_guard<std::mutex> lk(g_log_mu);
    if (g_file && g_file != stderr) fclose(g_file);
    g_file = nullptr;
    if (g_threshold.load() <= lsNONE) return;
    printf("log will go to: %s\n", path.c_str());
    g_file = fopen(path.c_str(), "a");
    if (!g_file) {
        fprintf(stderr, "log will go to stderr\n");
        g_file = stderr;
    }
}
}  // namespace

void log_set_threshold(int s) { g_threshold.store(s < lsNONE ? lsNONE : (s > lsTRACE ? lsTRACE : s)); }
int log_get_threshold() { return g_threshold.load(std::memory_order_relaxed); }
void log_set_to_unique_file(bool on) { g_unique_file.store(on); }
void log_set_host(const HostCallbacks* cb) {
    std::lock_guard<std::mutex> lk(g_log_mu);
    g_host = cb;
}

void log_open_netmerger(const std::string& dir) {
    if (!g_unique_file.load()) return;  // standard Hadoop logging configured: no own file
    open_file(dir + "udaNetMerger.log");
}

void log_open_mofsupplier(const std::string& dir) {
    if (!g_unique_file.load()) return;
    char host[100] = {0};
    if (gethostname(host, sizeof(host) - 1) != 0) snprintf(host, sizeof(host), "unknown");
    char user[100];
    struct passwd* pwd = getpwuid(getuid());
    if (pwd && pwd->pw_name) snprintf(user, sizeof(user), "%s", pwd->pw_name);
    else snprintf(user, sizeof(user), "uid-%u", (unsigned)getuid());
    open_file(dir + "/hadoop-" + user + "-udaMOFSupplier-" + host + ".log");
}

void log_close() {
    std::lock_guard<std::mutex> lk(g_log_mu);
    if (g_file && g_file != stderr) fclose(g_file);
    g_file = nullptr;
}

void log_write(const char* func, const char* file, int line, int severity, const char* fmt, ...) {
    if (severity <= lsNONE || severity > lsTRACE) return;
    char msg[1024];
    va_list ap;
    va_start(ap, fmt);
    int n = vsnprintf(msg, sizeof(msg), fmt, ap);
    va_end(ap);
    if (n < 0) return;
    const char* base = strrchr(file, '/');
    base = base ? base + 1 : file;
    if (!g_unique_file.load()) {
        // sink A: the host's logger; the location is appended to the message like the reference does
        size_t len = strnlen(msg, sizeof(msg));
        if (len < sizeof(msg) - 1) snprintf(msg + len, sizeof(msg) - len, " (%s:%d)", base, line);
        const HostCallbacks* h;
        {
            std::lock_guard<std::mutex> lk(g_log_mu);
            h = g_host;
        }
        if (h && h->log_to_java) h->log_to_java(h->user, msg, severity);
        else if (severity <= lsWARN) fprintf(stderr, "[uda %s] %s\n", kSeverityName[severity], msg);
        return;
    }
    time_t now = time(nullptr);
    struct tm tmv;
    localtime_r(&now, &tmv);
    std::lock_guard<std::mutex> lk(g_log_mu);
    FILE* f = g_file ? g_file : stderr;
    fprintf(f, "%02d:%02d:%02d %-5s [thr=%zx %s() %s:%d] %s\n", tmv.tm_hour, tmv.tm_min, tmv.tm_sec, kSeverityName[severity],
            std::hash<std::thread::id>()(std::this_thread::get_id()) & 0xffffff, func, base, line, msg);
    fflush(f);
}

std::string conf_get(const HostCallbacks* cb, const char* name, const char* dflt) {
    std::string v;
    if (cb && cb->get_conf_data && cb->get_conf_data(cb->user, name, dflt, &v)) return v;
    return dflt ? dflt : "";
}

}  // namespace uda
