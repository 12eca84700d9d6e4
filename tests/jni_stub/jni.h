// Minimal stand-in for the JDK's <jni.h>: only what java/jni/mp4j_jni.cc uses, with the JDK's signatures, so
// that the shim can be type-checked and linked on an image without a JDK (tests/test_jni_shim.py). It is NOT
// used to build the real library: java/build.sh compiles against $JAVA_HOME/include.
#pragma once
#include <stdint.h>

#define JNIEXPORT __attribute__((visibility("default")))
#define JNICALL
#define JNI_ABORT 2
#define JNI_COMMIT 1

typedef int32_t jint;
typedef int64_t jlong;
typedef int8_t jbyte;
typedef uint8_t jboolean;
typedef jint jsize;

// the "JVM" of the test harness (tests/cpp/jni_harness.cc): objects carry their payload directly
class _jobject {};
class _jclass : public _jobject {};
class _jstring : public _jobject {
 public:
  const char* utf = nullptr;
};
class _jarray : public _jobject {
 public:
  void* data = nullptr;  // elements
  jsize len = 0;         // element count
  int critical = 0;      // open Get/ReleasePrimitiveArrayCritical regions (the harness checks none is open while blocking)
};
class _jbyteArray : public _jarray {};
class _jlongArray : public _jarray {};
typedef _jobject* jobject;
typedef _jclass* jclass;
typedef _jstring* jstring;
typedef _jarray* jarray;
typedef _jbyteArray* jbyteArray;
typedef _jlongArray* jlongArray;

struct JNIEnv {
  jsize GetArrayLength(jarray a);
  jbyte* GetByteArrayElements(jbyteArray a, jboolean* is_copy);
  void ReleaseByteArrayElements(jbyteArray a, jbyte* p, jint mode);
  jlong* GetLongArrayElements(jlongArray a, jboolean* is_copy);
  void ReleaseLongArrayElements(jlongArray a, jlong* p, jint mode);
  void* GetPrimitiveArrayCritical(jarray a, jboolean* is_copy);
  void ReleasePrimitiveArrayCritical(jarray a, void* p, jint mode);
  const char* GetStringUTFChars(jstring s, jboolean* is_copy);
  void ReleaseStringUTFChars(jstring s, const char* c);
  jstring NewStringUTF(const char* c);
  jbyteArray NewByteArray(jsize n);
  void SetByteArrayRegion(jbyteArray a, jsize start, jsize n, const jbyte* buf);
};
