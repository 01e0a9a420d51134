// Mock-JVM test of the JNI bridge: a fake JavaVM / JNIEnv function table (same slot layout as a real JVM,
// see csrc/jni/jni.h) records every call the bridge makes, plays the Java side of the up-calls, and checks
// names, signatures, the failure contract and a full provider -> consumer data path through
// dataFromUda(DirectByteBuffer, len). Exit code 0 = all checks passed.
#include <dlfcn.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../csrc/jni/jni.h"
#include "../../include/uda_rt_api.h"

#define CHECK(c)                                                                 \
    do {                                                                         \
        if (!(c)) { fprintf(stderr, "CHECK failed %s:%d: %s\n", __FILE__, __LINE__, #c); exit(1); } \
    } while (0)

// ---- mock object model ----------------------------------------------------------------------------------
struct Obj {
    enum Kind { CLASS, STRING, ARRAY, BUFFER, RECORD, METHOD, FIELD } kind;
    std::string s;                 // class name / string value / method "name|sig" / field name
    std::vector<Obj*> elems;       // ARRAY
    void* addr = nullptr;          // BUFFER
    long long cap = 0;
    long long off = 0, raw = 0, part = 0;  // RECORD
    std::string path;
};
static std::mutex g_mu;
static std::vector<std::string> g_calls;
static std::vector<std::pair<int, std::string>> g_logs;
static std::string g_stream;
static std::atomic<int> g_fetch_over{0}, g_failures{0}, g_attached{0}, g_detached{0};
static std::string g_thrown_class, g_thrown_msg;
static std::map<std::string, Obj*> g_methods;
static std::map<std::string, std::pair<std::string, std::vector<long long>>> g_mofs;  // map id -> (path, off/raw/part per reduce)
static int g_direct_buffers = 0;
static bool g_buffer_cap_ok = true;   // every dataFromUda buffer had capacity == len

static Obj* mk(Obj::Kind k, const std::string& s = "") { auto* o = new Obj(); o->kind = k; o->s = s; return o; }
static void note(const std::string& s) { std::lock_guard<std::mutex> lk(g_mu); g_calls.push_back(s); }
static jobject J(Obj* o) { return (jobject)o; }
static Obj* O(jobject j) { return (Obj*)j; }

// ---- JNIEnv functions -------------------------------------------------------------------------------------
static jint m_GetVersion(JNIEnv*) { return JNI_VERSION_1_6; }
static jclass m_FindClass(JNIEnv*, const char* n) { note(std::string("FindClass ") + n); return J(mk(Obj::CLASS, n)); }
static jint m_ThrowNew(JNIEnv*, jclass c, const char* msg) { g_thrown_class = O(c)->s; g_thrown_msg = msg ? msg : ""; return 0; }
static jthrowable m_ExceptionOccurred(JNIEnv*) { return nullptr; }
static void m_ExceptionDescribe(JNIEnv*) {}
static void m_ExceptionClear(JNIEnv*) {}
static jobject m_NewGlobalRef(JNIEnv*, jobject o) { note("NewGlobalRef"); return o; }
static void m_DeleteGlobalRef(JNIEnv*, jobject) {}
static void m_DeleteLocalRef(JNIEnv*, jobject) {}
static jclass m_GetObjectClass(JNIEnv*, jobject) { return J(mk(Obj::CLASS, "org/apache/hadoop/mapred/IndexRecordBridge")); }
static jfieldID m_GetFieldID(JNIEnv*, jclass, const char* n, const char* sig) {
    note(std::string("GetFieldID ") + n + " " + sig);
    return (jfieldID)mk(Obj::FIELD, n);
}
static jobject m_GetObjectField(JNIEnv*, jobject o, jfieldID f) {
    CHECK(((Obj*)f)->s == "pathMOF");
    return J(mk(Obj::STRING, O(o)->path));
}
static jlong m_GetLongField(JNIEnv*, jobject o, jfieldID f) {
    const std::string& n = ((Obj*)f)->s;
    return n == "startOffset" ? O(o)->off : n == "rawLength" ? O(o)->raw : O(o)->part;
}
static jmethodID m_GetStaticMethodID(JNIEnv*, jclass, const char* n, const char* sig) {
    std::string key = std::string(n) + "|" + sig;
    note("GetStaticMethodID " + key);
    static const char* known[] = {"fetchOverMessage|()V", "dataFromUda|(Ljava/lang/Object;I)V",
                                  "getPathUda|(Ljava/lang/String;Ljava/lang/String;I)Ljava/lang/Object;",
                                  "getConfData|(Ljava/lang/String;Ljava/lang/String;)Ljava/lang/String;",
                                  "logToJava|(Ljava/lang/String;I)V", "failureInUda|()V"};
    for (auto k : known) if (key == k) { auto* m = mk(Obj::METHOD, key); g_methods[key] = m; return (jmethodID)m; }
    return nullptr;  // wrong name or signature: a real JVM would throw NoSuchMethodError
}
static jobject m_CallStaticObjectMethod(JNIEnv*, jclass, jmethodID mid, ...) {
    const std::string& m = ((Obj*)mid)->s;
    va_list ap;
    va_start(ap, mid);
    jobject ret = nullptr;
    if (m.rfind("getPathUda", 0) == 0) {
        Obj* job = O(va_arg(ap, jobject));
        Obj* map = O(va_arg(ap, jobject));
        int reduce = va_arg(ap, jint);
        note("getPathUda " + job->s + " " + map->s + " " + std::to_string(reduce));
        auto it = g_mofs.find(map->s);
        if (it != g_mofs.end()) {
            Obj* r = mk(Obj::RECORD);
            r->off = it->second.second[reduce * 3]; r->raw = it->second.second[reduce * 3 + 1]; r->part = it->second.second[reduce * 3 + 2];
            r->path = it->second.first;
            ret = J(r);
        }
    } else if (m.rfind("getConfData", 0) == 0) {
        Obj* name = O(va_arg(ap, jobject));
        Obj* dflt = O(va_arg(ap, jobject));
        (void)name;
        ret = J(mk(Obj::STRING, dflt->s));  // JobConf.get(name, default) with nothing configured
    }
    va_end(ap);
    return ret;
}
static void m_CallStaticVoidMethod(JNIEnv*, jclass, jmethodID mid, ...) {
    const std::string& m = ((Obj*)mid)->s;
    va_list ap;
    va_start(ap, mid);
    if (m.rfind("fetchOverMessage", 0) == 0) ++g_fetch_over;
    else if (m.rfind("failureInUda", 0) == 0) ++g_failures;
    else if (m.rfind("logToJava", 0) == 0) {
        Obj* s = O(va_arg(ap, jobject));
        int sev = va_arg(ap, jint);
        std::lock_guard<std::mutex> lk(g_mu);
        g_logs.emplace_back(sev, s->s);
    } else if (m.rfind("dataFromUda", 0) == 0) {
        Obj* b = O(va_arg(ap, jobject));
        int len = va_arg(ap, jint);
        CHECK(b->kind == Obj::BUFFER && len <= b->cap);
        if (b->cap != len) g_buffer_cap_ok = false;
        std::lock_guard<std::mutex> lk(g_mu);
        g_stream.append((const char*)b->addr, (size_t)len);  // what UdaPluginRT.dataFromUda does with directBuf.get()
    }
    va_end(ap);
}
static jstring m_NewStringUTF(JNIEnv*, const char* s) { return J(mk(Obj::STRING, s ? s : "")); }
static const char* m_GetStringUTFChars(JNIEnv*, jstring s, jboolean*) { return O(s)->s.c_str(); }
static void m_ReleaseStringUTFChars(JNIEnv*, jstring, const char*) {}
static jsize m_GetArrayLength(JNIEnv*, jarray a) { return (jsize)O(a)->elems.size(); }
static jobject m_GetObjectArrayElement(JNIEnv*, jobjectArray a, jsize i) { return J(O(a)->elems[(size_t)i]); }
static jweak m_NewWeakGlobalRef(JNIEnv*, jobject o) { note("NewWeakGlobalRef"); return o; }
static void m_DeleteWeakGlobalRef(JNIEnv*, jweak) {}
static jboolean m_ExceptionCheck(JNIEnv*) { return 0; }
static jobject m_NewDirectByteBuffer(JNIEnv*, void* addr, jlong cap) {
    ++g_direct_buffers;
    Obj* b = mk(Obj::BUFFER);
    b->addr = addr; b->cap = cap;
    return J(b);
}
static void* m_GetDirectBufferAddress(JNIEnv*, jobject b) { return O(b)->addr; }
static jlong m_GetDirectBufferCapacity(JNIEnv*, jobject b) { return O(b)->cap; }

static JNINativeInterface_ g_env_table;
static const JNINativeInterface_* g_env_ptr = &g_env_table;
static JNIInvokeInterface_ g_vm_table;
static const JNIInvokeInterface_* g_vm_ptr = &g_vm_table;
static thread_local bool t_attached = false;
static std::thread::id g_main_thread;

static jint v_GetEnv(JavaVM*, void** penv, jint) {
    if (std::this_thread::get_id() == g_main_thread || t_attached) { *penv = (void*)&g_env_ptr; return JNI_OK; }
    *penv = nullptr;
    return JNI_EDETACHED;
}
static jint v_Attach(JavaVM*, void** penv, void*) { t_attached = true; ++g_attached; *penv = (void*)&g_env_ptr; return JNI_OK; }
static jint v_Detach(JavaVM*) { t_attached = false; ++g_detached; return JNI_OK; }

// ---- helpers ------------------------------------------------------------------------------------------------
static int vint_put(std::string& s, long long v) {
    unsigned char b[9];
    int n = uda_rt_vint_encode(v, b);
    s.append((char*)b, (size_t)n);
    return n;
}
static std::string segment(const std::vector<std::pair<std::string, std::string>>& recs) {
    std::string s;
    for (auto& kv : recs) { vint_put(s, (long long)kv.first.size()); vint_put(s, (long long)kv.second.size()); s += kv.first; s += kv.second; }
    s += "\xff\xff";
    return s;
}
static Obj* jarray_of(const std::vector<std::string>& v) {
    Obj* a = mk(Obj::ARRAY);
    for (auto& s : v) a->elems.push_back(mk(Obj::STRING, s));
    return a;
}
static bool has_call(const std::string& s) {
    std::lock_guard<std::mutex> lk(g_mu);
    for (auto& c : g_calls) if (c == s) return true;
    return false;
}
static bool has_log(const std::string& needle) {
    std::lock_guard<std::mutex> lk(g_mu);
    for (auto& l : g_logs) if (l.second.find(needle) != std::string::npos) return true;
    return false;
}

typedef jint (*OnLoad)(JavaVM*, void*);
typedef void (*OnUnload)(JavaVM*, void*);
typedef jint (*StartNative)(JNIEnv*, jclass, jboolean, jobjectArray, jint, jboolean);
typedef void (*DoCommand)(JNIEnv*, jclass, jstring);
typedef void (*ReduceExit)(JNIEnv*, jclass);
typedef void (*SetLogLevel)(JNIEnv*, jclass, jint);

int main(int argc, char** argv) {
    CHECK(argc >= 3);
    void* lib = dlopen(argv[1], RTLD_NOW);
    if (!lib) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 2; }
    const std::string tmp = argv[2];
    g_main_thread = std::this_thread::get_id();

    memset(&g_env_table, 0, sizeof(g_env_table));
    g_env_table.GetVersion = m_GetVersion; g_env_table.FindClass = m_FindClass; g_env_table.ThrowNew = m_ThrowNew;
    g_env_table.ExceptionOccurred = m_ExceptionOccurred; g_env_table.ExceptionDescribe = m_ExceptionDescribe;
    g_env_table.ExceptionClear = m_ExceptionClear; g_env_table.NewGlobalRef = m_NewGlobalRef; g_env_table.DeleteGlobalRef = m_DeleteGlobalRef;
    g_env_table.DeleteLocalRef = m_DeleteLocalRef; g_env_table.GetObjectClass = m_GetObjectClass; g_env_table.GetFieldID = m_GetFieldID;
    g_env_table.GetObjectField = m_GetObjectField; g_env_table.GetLongField = m_GetLongField; g_env_table.GetStaticMethodID = m_GetStaticMethodID;
    g_env_table.CallStaticObjectMethod = m_CallStaticObjectMethod; g_env_table.CallStaticVoidMethod = m_CallStaticVoidMethod;
    g_env_table.NewStringUTF = m_NewStringUTF; g_env_table.GetStringUTFChars = m_GetStringUTFChars;
    g_env_table.ReleaseStringUTFChars = m_ReleaseStringUTFChars; g_env_table.GetArrayLength = m_GetArrayLength;
    g_env_table.GetObjectArrayElement = m_GetObjectArrayElement; g_env_table.NewWeakGlobalRef = m_NewWeakGlobalRef;
    g_env_table.DeleteWeakGlobalRef = m_DeleteWeakGlobalRef; g_env_table.ExceptionCheck = m_ExceptionCheck;
    g_env_table.NewDirectByteBuffer = m_NewDirectByteBuffer; g_env_table.GetDirectBufferAddress = m_GetDirectBufferAddress;
    g_env_table.GetDirectBufferCapacity = m_GetDirectBufferCapacity;
    memset(&g_vm_table, 0, sizeof(g_vm_table));
    g_vm_table.GetEnv = v_GetEnv; g_vm_table.AttachCurrentThread = v_Attach; g_vm_table.DetachCurrentThread = v_Detach;
    JavaVM* vm = (JavaVM*)&g_vm_ptr;
    JNIEnv* env = (JNIEnv*)&g_env_ptr;

    auto onload = (OnLoad)dlsym(lib, "JNI_OnLoad");
    auto onunload = (OnUnload)dlsym(lib, "JNI_OnUnload");
    auto start = (StartNative)dlsym(lib, "Java_com_mellanox_hadoop_mapred_UdaBridge_startNative");
    auto docmd = (DoCommand)dlsym(lib, "Java_com_mellanox_hadoop_mapred_UdaBridge_doCommandNative");
    auto rexit = (ReduceExit)dlsym(lib, "Java_com_mellanox_hadoop_mapred_UdaBridge_reduceExitMsgNative");
    auto setlog = (SetLogLevel)dlsym(lib, "Java_com_mellanox_hadoop_mapred_UdaBridge_setLogLevelNative");
    CHECK(onload && onunload && start && docmd && rexit && setlog);  // the exact exported symbol names

    // 1. JNI_OnLoad caches the class (weak global ref) and the five call-backs by name + signature
    CHECK(onload(vm, nullptr) == JNI_VERSION_1_4);
    CHECK(has_call("FindClass com/mellanox/hadoop/mapred/UdaBridge"));
    CHECK(has_call("NewWeakGlobalRef"));
    CHECK(has_call("GetStaticMethodID fetchOverMessage|()V"));
    CHECK(has_call("GetStaticMethodID dataFromUda|(Ljava/lang/Object;I)V"));
    CHECK(has_call("GetStaticMethodID getPathUda|(Ljava/lang/String;Ljava/lang/String;I)Ljava/lang/Object;"));
    CHECK(has_call("GetStaticMethodID getConfData|(Ljava/lang/String;Ljava/lang/String;)Ljava/lang/String;"));
    CHECK(has_call("GetStaticMethodID logToJava|(Ljava/lang/String;I)V"));
    CHECK(!has_call("GetStaticMethodID failureInUda|()V"));  // looked up lazily

    // 2. consumer: an unsupported key class in INIT becomes a Java UdaRuntimeException on the JNI thread
    jclass cls = J(mk(Obj::CLASS, "com/mellanox/hadoop/mapred/UdaBridge"));
    CHECK(start(env, cls, 1, J(jarray_of({"-w", "256", "-r", "1", "-a", "1", "-m", "1", "-g", tmp + "/", "-s", "64"})), 4, 0) == 0);
    CHECK(has_log("UDA version is "));
    docmd(env, cls, J(mk(Obj::STRING, "12:7:1:job_1:attempt_r_0:0:65536:4096:org.apache.hadoop.io.FloatWritable:null:0:1048576:0")));
    CHECK(g_thrown_class == "com/mellanox/hadoop/mapred/UdaRuntimeException");
    CHECK(g_thrown_msg.find("unsupported type") != std::string::npos);
    g_thrown_class.clear();
    docmd(env, cls, J(mk(Obj::STRING, "5:4:h:j:m:0")));  // ignored after the failure: no second exception
    CHECK(g_thrown_class.empty());
    rexit(env, cls);

    // 3. provider through JNI (getPathUda up-call + field reads) serving a consumer driven through the C API
    std::vector<std::pair<std::string, std::string>> r0, r1;
    for (int i = 0; i < 3000; ++i) {
        char k[16];
        snprintf(k, sizeof(k), "%08d", i * 2);
        r0.emplace_back(std::string(k, 8), std::string((size_t)(i % 50), 'a'));
        snprintf(k, sizeof(k), "%08d", i * 2 + 1);
        r1.emplace_back(std::string(k, 8), std::string((size_t)(i % 70), 'b'));
    }
    std::string expect;
    {
        std::vector<std::pair<std::string, std::string>> all;
        for (size_t i = 0; i < r0.size(); ++i) { all.push_back(r0[i]); all.push_back(r1[i]); }
        expect = segment(all);
    }
    int mi = 0;
    for (auto* recs : {&r0, &r1}) {
        std::string seg = segment(*recs) + std::string("\0\0\0\0", 4);
        std::string path = tmp + "/jni_map" + std::to_string(mi) + ".out";
        FILE* f = fopen(path.c_str(), "wb");
        CHECK(f && fwrite(seg.data(), 1, seg.size(), f) == seg.size());
        fclose(f);
        g_mofs["attempt_m_" + std::to_string(mi)] = {path, {0, (long long)seg.size() - 4, (long long)seg.size()}};
        ++mi;
    }
    CHECK(start(env, cls, 0, J(jarray_of({"-w", "256", "-r", "0", "-m", "1", "-g", tmp, "-s", "16"})), 4, 0) == 0);
    CHECK(has_log("The version is "));
    // find the provider's port from its log line
    int port = 0;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        for (auto& l : g_logs) {
            size_t p = l.second.find("MOFSupplier ready: port=");
            if (p != std::string::npos) port = atoi(l.second.c_str() + p + 24);
        }
    }
    CHECK(port > 0);
    struct Sink {
        std::mutex mu;
        std::string data;
        std::atomic<int> over{0}, fail{0};
        size_t size() { std::lock_guard<std::mutex> lk(mu); return data.size(); }
    } sink;
    uda_host_callbacks_t cb{};
    cb.user = &sink;
    cb.data_from_uda = [](void* u, const void* b, int n) {
        std::lock_guard<std::mutex> lk(((Sink*)u)->mu);
        ((Sink*)u)->data.append((const char*)b, (size_t)n);
    };
    cb.fetch_over_message = [](void* u) { ++((Sink*)u)->over; };
    cb.failure_in_uda = [](void* u) { ++((Sink*)u)->fail; };
    std::string ports = std::to_string(port);
    const char* cargv[] = {"-w", "256", "-r", ports.c_str(), "-a", "1", "-m", "1", "-g", "/tmp/", "-s", "16"};
    uda_role_t cons = uda_rt_start(1, 12, cargv, 4, 0, &cb);
    CHECK(!uda_rt_failed(cons));
    CHECK(uda_rt_do_command(cons, "12:7:2:job_1:attempt_r_0:0:16384:4096:org.apache.hadoop.io.LongWritable:null:0:67108864:0") == 0);
    CHECK(uda_rt_do_command(cons, "5:4:127.0.0.1:job_1:attempt_m_0:0") == 0);
    CHECK(uda_rt_do_command(cons, "5:4:127.0.0.1:job_1:attempt_m_1:0") == 0);
    for (int i = 0; i < 3000 && sink.size() < expect.size() && !sink.fail; ++i) std::this_thread::sleep_for(std::chrono::milliseconds(10));
    { std::lock_guard<std::mutex> lk(sink.mu); CHECK(sink.fail == 0 && sink.data == expect && sink.over >= 1); }
    CHECK(has_call("getPathUda job_1 attempt_m_0 0") && has_call("getPathUda job_1 attempt_m_1 0"));
    CHECK(has_call("GetFieldID startOffset J") && has_call("GetFieldID rawLength J") && has_call("GetFieldID partLength J") &&
          has_call("GetFieldID pathMOF Ljava/lang/String;"));
    CHECK(g_attached >= 1);  // the provider's engine thread attached itself for the up-call
    uda_rt_reduce_exit(cons);
    uda_rt_destroy(cons);
    docmd(env, cls, J(mk(Obj::STRING, "1:0")));  // EXIT to the provider

    // 4. consumer through JNI fed by a provider driven through the C API: dataFromUda(DirectByteBuffer, len)
    struct Prov { } pv;
    uda_host_callbacks_t pcb{};
    pcb.user = &pv;
    pcb.get_path_uda = [](void*, const char*, const char* map, int reduce, long long* off, long long* raw, long long* part, char* out, int cap) -> int {
        auto it = g_mofs.find(map);
        if (it == g_mofs.end()) return 0;
        *off = it->second.second[reduce * 3]; *raw = it->second.second[reduce * 3 + 1]; *part = it->second.second[reduce * 3 + 2];
        snprintf(out, (size_t)cap, "%s", it->second.first.c_str());
        return 1;
    };
    const char* pargv[] = {"-w", "256", "-r", "0", "-m", "1", "-g", "/tmp", "-s", "16"};
    uda_role_t prov = uda_rt_start(0, 10, pargv, 4, 0, &pcb);
    CHECK(!uda_rt_failed(prov) && uda_rt_port(prov) > 0);
    ports = std::to_string(uda_rt_port(prov));
    g_stream.clear();
    g_fetch_over = 0;
    // (optional third program argument: the merge approach of this consumer; "3" = device merge, which the test driver
    //  runs against the host-only mock of the kernel library: blocks then come out of the pinned delivery windows at
    //  arbitrary offsets, every one through a direct buffer of exactly its length)
    const std::string approach = argc >= 4 ? argv[3] : "1";
    CHECK(start(env, cls, 1, J(jarray_of({"-w", "256", "-r", ports, "-a", approach, "-m", "1", "-g", "/tmp/", "-s", "16"})), 4, 0) == 0);
    docmd(env, cls, J(mk(Obj::STRING, "12:7:2:job_1:attempt_r_0:0:16384:4096:org.apache.hadoop.io.LongWritable:null:0:67108864:0")));
    docmd(env, cls, J(mk(Obj::STRING, "5:4:127.0.0.1:job_1:attempt_m_0:0")));
    docmd(env, cls, J(mk(Obj::STRING, "5:4:127.0.0.1:job_1:attempt_m_1:0")));
    for (int i = 0; i < 3000; ++i) {
        { std::lock_guard<std::mutex> lk(g_mu); if (g_stream.size() >= expect.size()) break; }
        std::this_thread::sleep_for(std::chrono::milliseconds(10));
    }
    { std::lock_guard<std::mutex> lk(g_mu); CHECK(g_stream == expect); }
    CHECK(g_fetch_over >= 1 && g_failures == 0);
    CHECK(g_direct_buffers >= 1 && g_buffer_cap_ok);  // one buffer per delivered block, capacity == delivered length
    if (approach == "3") CHECK(has_log("device merge of 2 segments"));   // the device path really served this stream
    rexit(env, cls);

    // 5. consumer whose provider is gone: the merge thread's failure arrives as failureInUda()
    uda_rt_destroy(prov);
    CHECK(start(env, cls, 1, J(jarray_of({"-w", "256", "-r", "1", "-a", "1", "-m", "1", "-g", "/tmp/", "-s", "16"})), 4, 0) == 0);
    docmd(env, cls, J(mk(Obj::STRING, "12:7:1:job_1:attempt_r_0:0:16384:4096:org.apache.hadoop.io.LongWritable:null:0:67108864:0")));
    docmd(env, cls, J(mk(Obj::STRING, "5:4:127.0.0.1:job_1:attempt_m_0:0")));
    for (int i = 0; i < 2000 && g_failures == 0; ++i) std::this_thread::sleep_for(std::chrono::milliseconds(10));
    CHECK(g_failures == 1);
    CHECK(has_call("GetStaticMethodID failureInUda|()V"));
    rexit(env, cls);
    setlog(env, cls, 6);

    onunload(vm, nullptr);
    CHECK(g_detached >= 1);  // native threads detached when they ended
    printf("jni mock: all checks passed (%zu recorded calls, %d attaches, %d detaches)\n", g_calls.size(), g_attached.load(), g_detached.load());
    return 0;
}
