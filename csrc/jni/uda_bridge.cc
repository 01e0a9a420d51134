// JNI bridge: the native half of com.mellanox.hadoop.mapred.UdaBridge.
//
// Same exported symbols, method names and JNI signatures as the reference (src/UdaBridge.cc):
//   JNI_OnLoad / JNI_OnUnload                                    :110-185
//   Java_com_mellanox_hadoop_mapred_UdaBridge_startNative        :187-263  (Z[Ljava/lang/String;IZ)I
//   ..._doCommandNative                                          :266-295  (Ljava/lang/String;)V
//   ..._reduceExitMsgNative                                      :299-316  ()V
//   ..._setLogLevelNative                                        :318-330  (I)V
// and the up-calls looked up by name + signature in JNI_OnLoad (:138-171):
//   fetchOverMessage()V, dataFromUda(Ljava/lang/Object;I)V,
//   getPathUda(Ljava/lang/String;Ljava/lang/String;I)Ljava/lang/Object;  (fields startOffset,rawLength,partLength:J, pathMOF:String)
//   getConfData(Ljava/lang/String;Ljava/lang/String;)Ljava/lang/String;, logToJava(Ljava/lang/String;I)V,
//   failureInUda()V (looked up lazily, :516).
// Failure contract (:81-107, :506-530): an exception on a JNI thread of the consumer becomes a Java
// UdaRuntimeException; a failure on a native thread becomes a failureInUda() up-call; afterwards
// commands are ignored. Native threads attach to the JVM on their first up-call and detach when they
// end. The role logic itself lives behind the C API (include/uda_rt_api.h).
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/uda_rt_api.h"
#include "jni.h"

namespace {

JavaVM* g_vm = nullptr;
jweak g_cls_weak = nullptr;   // weak global ref: lets the class (and this library) be unloaded
jclass g_cls = nullptr;
jmethodID g_fetchOver, g_dataFromUda, g_getPathUda, g_getConfData, g_logToJava;
jfieldID g_fidOffset, g_fidRaw, g_fidPart, g_fidPath;
uda_role_t g_role = nullptr;
bool g_is_net_merger = false;

#define ENV(fn, ...) ((*env)->fn(env, ##__VA_ARGS__))

// Detaches a native thread from the JVM when it ends (reference: uda_thread_create wraps every thread
// with attach/detach, src/CommUtils/UdaUtil.cc:39-94).
struct ThreadAttachment {
    bool attached = false;
    ~ThreadAttachment() {
        if (attached && g_vm) (*g_vm)->DetachCurrentThread(g_vm);
    }
};
thread_local ThreadAttachment t_attachment;

JNIEnv* thread_env() {
    if (!g_vm) return nullptr;
    JNIEnv* env = nullptr;
    jint rc = (*g_vm)->GetEnv(g_vm, (void**)&env, JNI_VERSION_1_4);
    if (rc == JNI_OK) return env;
    if ((*g_vm)->AttachCurrentThread(g_vm, (void**)&env, nullptr) < 0) return nullptr;
    t_attachment.attached = true;
    return env;
}

// ---- up-call trampolines (uda_host_callbacks_t) --------------------------------------------------------
void up_fetch_over(void*) {
    JNIEnv* env = thread_env();
    if (env) ENV(CallStaticVoidMethod, g_cls, g_fetchOver);
}

// One block of whole records for Java. The block may live anywhere (the two 1 MiB staging blocks of the CPU merges, an
// arbitrary offset inside the pinned windows of the device merge, a temporary for a record larger than a block), so the
// DirectByteBuffer is made for exactly this call — capacity = len, never more than the memory behind it — and dropped
// afterwards. Nothing is cached by address: a cache keyed on the block address leaks one global reference per distinct
// address and keeps references to freed temporaries.
void up_data(void*, const void* buf, int len) {
    JNIEnv* env = thread_env();
    if (!env || len < 0) return;
    jobject jbuf = ENV(NewDirectByteBuffer, const_cast<void*>(buf), (jlong)len);
    if (!jbuf) return;
    ENV(CallStaticVoidMethod, g_cls, g_dataFromUda, jbuf, (jint)len);
    ENV(DeleteLocalRef, jbuf);
}

int up_get_path(void*, const char* job, const char* map, int reduce, long long* off, long long* raw, long long* part,
                char* path_out, int cap) {
    JNIEnv* env = thread_env();
    if (!env) return 0;
    jstring jjob = ENV(NewStringUTF, job), jmap = ENV(NewStringUTF, map);
    jobject rec = ENV(CallStaticObjectMethod, g_cls, g_getPathUda, jjob, jmap, (jint)reduce);
    ENV(DeleteLocalRef, jjob);
    ENV(DeleteLocalRef, jmap);
    if (!rec) return 0;
    if (!g_fidOffset) {
        jclass rc = ENV(GetObjectClass, rec);
        g_fidOffset = ENV(GetFieldID, rc, "startOffset", "J");
        g_fidRaw = ENV(GetFieldID, rc, "rawLength", "J");
        g_fidPart = ENV(GetFieldID, rc, "partLength", "J");
        g_fidPath = ENV(GetFieldID, rc, "pathMOF", "Ljava/lang/String;");
        if (!g_fidOffset || !g_fidRaw || !g_fidPart || !g_fidPath) { g_fidOffset = nullptr; return 0; }
    }
    *off = ENV(GetLongField, rec, g_fidOffset);
    *raw = ENV(GetLongField, rec, g_fidRaw);
    *part = ENV(GetLongField, rec, g_fidPart);
    jstring jpath = (jstring)ENV(GetObjectField, rec, g_fidPath);
    if (!jpath) return 0;
    const char* s = ENV(GetStringUTFChars, jpath, nullptr);
    snprintf(path_out, (size_t)cap, "%s", s ? s : "");
    if (s) ENV(ReleaseStringUTFChars, jpath, s);
    ENV(DeleteLocalRef, jpath);
    ENV(DeleteLocalRef, rec);
    return 1;
}

int up_get_conf(void*, const char* name, const char* dflt, char* out, int cap) {
    JNIEnv* env = thread_env();
    if (!env) return 0;
    jstring jn = ENV(NewStringUTF, name), jd = ENV(NewStringUTF, dflt);
    jstring jv = (jstring)ENV(CallStaticObjectMethod, g_cls, g_getConfData, jn, jd);
    ENV(DeleteLocalRef, jn);
    ENV(DeleteLocalRef, jd);
    if (!jv) return 0;
    const char* s = ENV(GetStringUTFChars, jv, nullptr);
    snprintf(out, (size_t)cap, "%s", s ? s : "");
    if (s) ENV(ReleaseStringUTFChars, jv, s);
    ENV(DeleteLocalRef, jv);
    return 1;
}

void up_log(void*, const char* msg, int severity) {
    JNIEnv* env = thread_env();
    if (!env) { fprintf(stderr, "[uda] %s\n", msg); return; }
    jstring jm = ENV(NewStringUTF, msg);
    ENV(CallStaticVoidMethod, g_cls, g_logToJava, jm, (jint)severity);
    ENV(DeleteLocalRef, jm);
}

void up_failure(void*) {
    JNIEnv* env = thread_env();
    if (!env || !g_is_net_merger) return;
    jmethodID mid = ENV(GetStaticMethodID, g_cls, "failureInUda", "()V");  // looked up lazily, like the reference
    if (mid) ENV(CallStaticVoidMethod, g_cls, mid);
}

// a failure detected on the calling (JNI) thread: raise UdaRuntimeException in Java (consumer only)
void raise_in_jni_thread(JNIEnv* env, const char* msg) {
    if (!g_is_net_merger) return;
    jclass ex = ENV(FindClass, "com/mellanox/hadoop/mapred/UdaRuntimeException");
    if (!ex) return;
    ENV(ThrowNew, ex, msg);
    ENV(DeleteLocalRef, ex);
}

}  // namespace

extern "C" {

JNIEXPORT jint JNICALL JNI_OnLoad(JavaVM* vm, void*) {
    printf("-->> In C++ JNI_OnLoad\n");
    g_vm = vm;
    JNIEnv* env = nullptr;
    if ((*vm)->GetEnv(vm, (void**)&env, JNI_VERSION_1_4) != JNI_OK) return JNI_ERR;  // direct buffers need 1.4
    jclass cls = ENV(FindClass, "com/mellanox/hadoop/mapred/UdaBridge");
    if (!cls) return JNI_ERR;
    g_cls_weak = ENV(NewWeakGlobalRef, cls);
    if (!g_cls_weak) return JNI_ERR;
    g_cls = (jclass)g_cls_weak;
    g_fetchOver = ENV(GetStaticMethodID, g_cls, "fetchOverMessage", "()V");
    g_dataFromUda = ENV(GetStaticMethodID, g_cls, "dataFromUda", "(Ljava/lang/Object;I)V");
    g_getPathUda = ENV(GetStaticMethodID, g_cls, "getPathUda", "(Ljava/lang/String;Ljava/lang/String;I)Ljava/lang/Object;");
    g_getConfData = ENV(GetStaticMethodID, g_cls, "getConfData", "(Ljava/lang/String;Ljava/lang/String;)Ljava/lang/String;");
    g_logToJava = ENV(GetStaticMethodID, g_cls, "logToJava", "(Ljava/lang/String;I)V");
    if (!g_fetchOver || !g_dataFromUda || !g_getPathUda || !g_getConfData || !g_logToJava) {
        printf("-->> In C++ a java UdaBridge callback method was NOT found\n");
        return JNI_ERR;
    }
    printf("-->> In C++ java callback methods were found and cached\n");
    return JNI_VERSION_1_4;
}

JNIEXPORT void JNICALL JNI_OnUnload(JavaVM* vm, void*) {
    JNIEnv* env = nullptr;
    if (g_role) { uda_rt_destroy(g_role); g_role = nullptr; }
    if (vm && (*vm)->GetEnv(vm, (void**)&env, JNI_VERSION_1_4) == JNI_OK) {
        if (g_cls_weak) ENV(DeleteWeakGlobalRef, g_cls_weak);
    }
    g_cls_weak = nullptr;
    g_cls = nullptr;
    g_vm = nullptr;
}

JNIEXPORT jint JNICALL Java_com_mellanox_hadoop_mapred_UdaBridge_startNative(JNIEnv* env, jclass, jboolean isNetMerger,
                                                                             jobjectArray args, jint log_level,
                                                                             jboolean log_to_uda_file) {
    const int argc = ENV(GetArrayLength, args);
    std::vector<std::string> store((size_t)argc);
    std::vector<const char*> argv((size_t)argc);
    for (int i = 0; i < argc; ++i) {
        jstring s = (jstring)ENV(GetObjectArrayElement, args, i);
        if (s) {
            const char* c = ENV(GetStringUTFChars, s, nullptr);
            store[i] = c ? c : "";
            if (c) ENV(ReleaseStringUTFChars, s, c);
            ENV(DeleteLocalRef, s);
        }
        argv[i] = store[i].c_str();
    }
    g_is_net_merger = isNetMerger != 0;
    uda_host_callbacks_t cb{};
    cb.fetch_over_message = up_fetch_over;
    cb.data_from_uda = up_data;
    cb.get_path_uda = up_get_path;
    cb.get_conf_data = up_get_conf;
    cb.log_to_java = up_log;
    cb.failure_in_uda = up_failure;
    if (g_role) { uda_rt_destroy(g_role); g_role = nullptr; }
    g_role = uda_rt_start(g_is_net_merger, argc, argv.data(), log_level, log_to_uda_file, &cb);
    if (uda_rt_failed(g_role)) {
        raise_in_jni_thread(env, uda_rt_last_error(g_role));
        return 1;
    }
    return 0;
}

JNIEXPORT void JNICALL Java_com_mellanox_hadoop_mapred_UdaBridge_doCommandNative(JNIEnv* env, jclass, jstring s) {
    if (!g_role) return;
    const char* c = ENV(GetStringUTFChars, s, nullptr);
    if (!c) { raise_in_jni_thread(env, "Out of Memory"); return; }
    std::string msg(c);
    ENV(ReleaseStringUTFChars, s, c);
    if (uda_rt_do_command(g_role, msg.c_str()) != 0) raise_in_jni_thread(env, uda_rt_last_error(g_role));
}

JNIEXPORT void JNICALL Java_com_mellanox_hadoop_mapred_UdaBridge_reduceExitMsgNative(JNIEnv*, jclass) {
    if (g_role) uda_rt_reduce_exit(g_role);
}

JNIEXPORT void JNICALL Java_com_mellanox_hadoop_mapred_UdaBridge_setLogLevelNative(JNIEnv*, jclass, jint level) {
    uda_rt_set_log_level(level);
}

}  // extern "C"
