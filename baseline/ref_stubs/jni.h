/* Minimal C++ view of <jni.h> for compiling the reference's Merger sources without a JDK (reference arm of
 * bench.py). Only what those sources name: the handle types and JNIEnv::DeleteWeakGlobalRef
 * (src/Merger/MergeManager.cc:186). The "JVM" behind it is the stub in baseline/ref_harness.cc; no Java runs. */
#ifndef UDA_REF_STUB_JNI_H
#define UDA_REF_STUB_JNI_H
#include <stdint.h>
#include <stddef.h>
#include <stdio.h>
#include <stdarg.h>
#include <pthread.h>   /* libstdc++ of the reference's era exposed pthread through <string>; today's does not */
typedef int32_t jint; typedef int64_t jlong; typedef int8_t jbyte; typedef uint8_t jboolean;
class _jobject {};
typedef _jobject* jobject; typedef jobject jclass; typedef jobject jstring; typedef jobject jweak; typedef jobject jthrowable;
struct _jmethodID; typedef _jmethodID* jmethodID;
struct JNIEnv_ {
    void DeleteWeakGlobalRef(jweak) {}
};
typedef JNIEnv_ JNIEnv;
struct JavaVM_ {}; typedef JavaVM_ JavaVM;
#define JNIEXPORT
#define JNICALL
#define JNI_OK 0
#endif
