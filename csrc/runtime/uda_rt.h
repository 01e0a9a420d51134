// uda_b200 host runtime — common definitions (errors, logging, host call-backs, version).
//
// Role in the system: this is the CPU side of the plug-in surface. It owns the text
// protocols, the MOF store, the TCP loop-back transport used for plumbing and the CPU merger; the GPU
// path is driven through libuda_kernels.so. Both the JNI bridge (csrc/jni) and the plain C API used by
// the Python "fake Hadoop" driver (c_api.cc) sit on the same core and reach the embedding application
// only through the HostCallbacks table below.
//
// Reference behaviour covered here: log facility with 7 severities and two sinks
// (src/include/IOUtility.h:151-165, src/CommUtils/IOUtility.cc:406-466,514-559) and UdaException with
// a backtrace captured at construction (src/CommUtils/IOUtility.cc:479-498,562-569).
#pragma once
#include <cstdarg>
#include <cstdint>
#include <stdexcept>
#include <string>

#ifndef VERSION_UDA
#define VERSION_UDA "b200-dev"
#endif

namespace uda {

// same numbering as the reference's log_severity_t: Java pushes these integers (UdaPlugin.java:80-91)
enum Severity { lsNONE = 0, lsFATAL = 1, lsERROR = 2, lsWARN = 3, lsINFO = 4, lsDEBUG = 5, lsTRACE = 6 };

class UdaError : public std::runtime_error {
public:
    explicit UdaError(const std::string& info);
    const std::string& info() const { return info_; }
    const std::string& full_message() const { return full_; }  // info + backtrace
private:
    std::string info_, full_;
};

std::string capture_backtrace(int skip = 1);

// One index-file entry as returned by the host's getPathUda up-call
// (reference: src/MOFServer/IndexInfo.h:98-104, fields read at src/UdaBridge.cc:352-415).
struct IndexRecord {
    int64_t offset = 0;       // startOffset inside file.out
    int64_t raw_length = 0;   // decompressed length incl. EOF marker
    int64_t part_length = 0;  // on-disk length
    std::string path;         // file.out path
};

// Up-calls into the embedding application (the JVM through JNI, or Python through ctypes).
// Signatures mirror UdaCallable / UdaBridge.java:85-145.
struct HostCallbacks {
    void* user = nullptr;
    void (*fetch_over_message)(void* user) = nullptr;
    void (*data_from_uda)(void* user, const void* buf, int len) = nullptr;
    // returns false when the host has no such map output
    bool (*get_path_uda)(void* user, const char* job, const char* map, int reduce, IndexRecord* out) = nullptr;
    // returns false when unavailable; value written to out
    bool (*get_conf_data)(void* user, const char* name, const char* dflt, std::string* out) = nullptr;
    void (*log_to_java)(void* user, const char* msg, int severity) = nullptr;
    void (*failure_in_uda)(void* user) = nullptr;
};

// ---- logging -----------------------------------------------------------------------------------------
void log_set_threshold(int severity);
int log_get_threshold();
void log_set_to_unique_file(bool on);              // false: route through HostCallbacks::log_to_java
void log_set_host(const HostCallbacks* cb);        // sink A
void log_open_netmerger(const std::string& dir);   // sink B: <dir>udaNetMerger.log
void log_open_mofsupplier(const std::string& dir); // sink B: <dir>/hadoop-<user>-udaMOFSupplier-<host>.log
void log_close();
void log_write(const char* func, const char* file, int line, int severity, const char* fmt, ...)
    __attribute__((format(printf, 5, 6)));

#define UDA_LOG(sev, ...)                                                                \
    do {                                                                                 \
        if ((sev) <= ::uda::log_get_threshold())                                         \
            ::uda::log_write(__func__, __FILE__, __LINE__, (sev), __VA_ARGS__);          \
    } while (0)

// getConfData with a default when the host has no answer (reference: src/UdaBridge.cc:419-438)
std::string conf_get(const HostCallbacks* cb, const char* name, const char* dflt);

}  // namespace uda
