/* Minimal clean-room subset of the Java Native Interface (JNI 1.6 layout), written from the public JNI
 * specification because no JDK exists on the build or GPU boxes. Only the slots the
 * bridge uses are named; every other slot is opaque padding so that the function-table offsets are those
 * of a real JVM. The static_asserts pin the indices the specification assigns (FindClass 6,
 * GetStaticMethodID 113, CallStaticVoidMethod 141, NewStringUTF 167, NewDirectByteBuffer 229, ...).
 *
 * The reference compiles against the JDK's jni.h (src/UdaBridge.cc:20-38). */
#ifndef UDA_VENDORED_JNI_H
#define UDA_VENDORED_JNI_H

#include <stdarg.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef uint8_t jboolean;
typedef int8_t jbyte;
typedef uint16_t jchar;
typedef int16_t jshort;
typedef int32_t jint;
typedef int64_t jlong;
typedef float jfloat;
typedef double jdouble;
typedef jint jsize;

struct _jobject;
typedef struct _jobject* jobject;
typedef jobject jclass;
typedef jobject jthrowable;
typedef jobject jstring;
typedef jobject jarray;
typedef jarray jobjectArray;
typedef jobject jweak;
struct _jfieldID;
typedef struct _jfieldID* jfieldID;
struct _jmethodID;
typedef struct _jmethodID* jmethodID;

#define JNI_FALSE 0
#define JNI_TRUE 1
#define JNI_OK 0
#define JNI_ERR (-1)
#define JNI_EDETACHED (-2)
#define JNI_EVERSION (-3)
#define JNI_VERSION_1_4 0x00010004
#define JNI_VERSION_1_6 0x00010006
#define JNIEXPORT __attribute__((visibility("default")))
#define JNIIMPORT
#define JNICALL

struct JNINativeInterface_;
struct JNIInvokeInterface_;
typedef const struct JNINativeInterface_* JNIEnv;   /* C view: JNIEnv* points at the table pointer */
typedef const struct JNIInvokeInterface_* JavaVM;

struct JNINativeInterface_ {
    void* unused_0_3[4];
    jint (*GetVersion)(JNIEnv*);  /* slot 4 */
    void* unused_5_5[1];
    jclass (*FindClass)(JNIEnv*, const char*);  /* slot 6 */
    void* unused_7_13[7];
    jint (*ThrowNew)(JNIEnv*, jclass, const char*);  /* slot 14 */
    jthrowable (*ExceptionOccurred)(JNIEnv*);  /* slot 15 */
    void (*ExceptionDescribe)(JNIEnv*);  /* slot 16 */
    void (*ExceptionClear)(JNIEnv*);  /* slot 17 */
    void* unused_18_20[3];
    jobject (*NewGlobalRef)(JNIEnv*, jobject);  /* slot 21 */
    void (*DeleteGlobalRef)(JNIEnv*, jobject);  /* slot 22 */
    void (*DeleteLocalRef)(JNIEnv*, jobject);  /* slot 23 */
    void* unused_24_30[7];
    jclass (*GetObjectClass)(JNIEnv*, jobject);  /* slot 31 */
    void* unused_32_93[62];
    jfieldID (*GetFieldID)(JNIEnv*, jclass, const char*, const char*);  /* slot 94 */
    jobject (*GetObjectField)(JNIEnv*, jobject, jfieldID);  /* slot 95 */
    void* unused_96_100[5];
    jlong (*GetLongField)(JNIEnv*, jobject, jfieldID);  /* slot 101 */
    void* unused_102_112[11];
    jmethodID (*GetStaticMethodID)(JNIEnv*, jclass, const char*, const char*);  /* slot 113 */
    jobject (*CallStaticObjectMethod)(JNIEnv*, jclass, jmethodID, ...);  /* slot 114 */
    void* unused_115_140[26];
    void (*CallStaticVoidMethod)(JNIEnv*, jclass, jmethodID, ...);  /* slot 141 */
    void* unused_142_166[25];
    jstring (*NewStringUTF)(JNIEnv*, const char*);  /* slot 167 */
    void* unused_168_168[1];
    const char* (*GetStringUTFChars)(JNIEnv*, jstring, jboolean*);  /* slot 169 */
    void (*ReleaseStringUTFChars)(JNIEnv*, jstring, const char*);  /* slot 170 */
    jsize (*GetArrayLength)(JNIEnv*, jarray);  /* slot 171 */
    void* unused_172_172[1];
    jobject (*GetObjectArrayElement)(JNIEnv*, jobjectArray, jsize);  /* slot 173 */
    void* unused_174_225[52];
    jweak (*NewWeakGlobalRef)(JNIEnv*, jobject);  /* slot 226 */
    void (*DeleteWeakGlobalRef)(JNIEnv*, jweak);  /* slot 227 */
    jboolean (*ExceptionCheck)(JNIEnv*);  /* slot 228 */
    jobject (*NewDirectByteBuffer)(JNIEnv*, void*, jlong);  /* slot 229 */
    void* (*GetDirectBufferAddress)(JNIEnv*, jobject);  /* slot 230 */
    jlong (*GetDirectBufferCapacity)(JNIEnv*, jobject);  /* slot 231 */
    void* unused_232_232[1];
};

struct JNIInvokeInterface_ {
    void* reserved0;
    void* reserved1;
    void* reserved2;
    jint (*DestroyJavaVM)(JavaVM*);
    jint (*AttachCurrentThread)(JavaVM*, void** penv, void* args);
    jint (*DetachCurrentThread)(JavaVM*);
    jint (*GetEnv)(JavaVM*, void** penv, jint version);
    jint (*AttachCurrentThreadAsDaemon)(JavaVM*, void** penv, void* args);
};

#ifdef __cplusplus
}
static_assert(offsetof(JNINativeInterface_, GetVersion) == 4 * sizeof(void*), "JNI slot 4");
static_assert(offsetof(JNINativeInterface_, FindClass) == 6 * sizeof(void*), "JNI slot 6");
static_assert(offsetof(JNINativeInterface_, ThrowNew) == 14 * sizeof(void*), "JNI slot 14");
static_assert(offsetof(JNINativeInterface_, ExceptionOccurred) == 15 * sizeof(void*), "JNI slot 15");
static_assert(offsetof(JNINativeInterface_, ExceptionDescribe) == 16 * sizeof(void*), "JNI slot 16");
static_assert(offsetof(JNINativeInterface_, ExceptionClear) == 17 * sizeof(void*), "JNI slot 17");
static_assert(offsetof(JNINativeInterface_, NewGlobalRef) == 21 * sizeof(void*), "JNI slot 21");
static_assert(offsetof(JNINativeInterface_, DeleteGlobalRef) == 22 * sizeof(void*), "JNI slot 22");
static_assert(offsetof(JNINativeInterface_, DeleteLocalRef) == 23 * sizeof(void*), "JNI slot 23");
static_assert(offsetof(JNINativeInterface_, GetObjectClass) == 31 * sizeof(void*), "JNI slot 31");
static_assert(offsetof(JNINativeInterface_, GetFieldID) == 94 * sizeof(void*), "JNI slot 94");
static_assert(offsetof(JNINativeInterface_, GetObjectField) == 95 * sizeof(void*), "JNI slot 95");
static_assert(offsetof(JNINativeInterface_, GetLongField) == 101 * sizeof(void*), "JNI slot 101");
static_assert(offsetof(JNINativeInterface_, GetStaticMethodID) == 113 * sizeof(void*), "JNI slot 113");
static_assert(offsetof(JNINativeInterface_, CallStaticObjectMethod) == 114 * sizeof(void*), "JNI slot 114");
static_assert(offsetof(JNINativeInterface_, CallStaticVoidMethod) == 141 * sizeof(void*), "JNI slot 141");
static_assert(offsetof(JNINativeInterface_, NewStringUTF) == 167 * sizeof(void*), "JNI slot 167");
static_assert(offsetof(JNINativeInterface_, GetStringUTFChars) == 169 * sizeof(void*), "JNI slot 169");
static_assert(offsetof(JNINativeInterface_, ReleaseStringUTFChars) == 170 * sizeof(void*), "JNI slot 170");
static_assert(offsetof(JNINativeInterface_, GetArrayLength) == 171 * sizeof(void*), "JNI slot 171");
static_assert(offsetof(JNINativeInterface_, GetObjectArrayElement) == 173 * sizeof(void*), "JNI slot 173");
static_assert(offsetof(JNINativeInterface_, NewWeakGlobalRef) == 226 * sizeof(void*), "JNI slot 226");
static_assert(offsetof(JNINativeInterface_, DeleteWeakGlobalRef) == 227 * sizeof(void*), "JNI slot 227");
static_assert(offsetof(JNINativeInterface_, ExceptionCheck) == 228 * sizeof(void*), "JNI slot 228");
static_assert(offsetof(JNINativeInterface_, NewDirectByteBuffer) == 229 * sizeof(void*), "JNI slot 229");
static_assert(offsetof(JNINativeInterface_, GetDirectBufferAddress) == 230 * sizeof(void*), "JNI slot 230");
static_assert(offsetof(JNINativeInterface_, GetDirectBufferCapacity) == 231 * sizeof(void*), "JNI slot 231");
static_assert(sizeof(JNINativeInterface_) == 233 * sizeof(void*), "JNI table size");
static_assert(offsetof(JNIInvokeInterface_, GetEnv) == 6 * sizeof(void*), "JavaVM slot 6");
#endif

#endif /* UDA_VENDORED_JNI_H */
