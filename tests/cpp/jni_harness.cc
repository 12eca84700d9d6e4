// Runs the JNI shim (java/jni/mp4j_jni.cc) without a JVM: this file implements the handful of JNIEnv functions the
// shim uses over plain C++ objects (tests/jni_stub/jni.h), then drives the shim's entry points the way
// com.fenbi.mp4j.comm.Native would: a master, p slaves (threads), host collectives on "double[]" / "float[]"
// arrays, byte transport. Checked: results, that the Java array is updated, that no critical region is open while a
// collective blocks, and that errors surface through lastError().
#include <jni.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <string>
#include <thread>
#include <vector>

#include "../../include/mp4j_types.h"

static std::atomic<int> g_fail{0};
static std::atomic<int> g_open_critical{0};
#define CHECK(c, ...)                      \
  do {                                     \
    if (!(c)) {                            \
      fprintf(stderr, "FAIL %s:%d: ", __FILE__, __LINE__); \
      fprintf(stderr, __VA_ARGS__);        \
      fprintf(stderr, "\n");               \
      g_fail.fetch_add(1);                 \
    }                                      \
  } while (0)

// ---- the fake JVM
jsize JNIEnv::GetArrayLength(jarray a) { return a->len; }
jbyte* JNIEnv::GetByteArrayElements(jbyteArray a, jboolean*) { return (jbyte*)a->data; }
void JNIEnv::ReleaseByteArrayElements(jbyteArray, jbyte*, jint) {}
jlong* JNIEnv::GetLongArrayElements(jlongArray a, jboolean*) { return (jlong*)a->data; }
void JNIEnv::ReleaseLongArrayElements(jlongArray, jlong*, jint) {}
void* JNIEnv::GetPrimitiveArrayCritical(jarray a, jboolean*) {
  a->critical++;
  g_open_critical.fetch_add(1);
  return a->data;
}
void JNIEnv::ReleasePrimitiveArrayCritical(jarray a, void*, jint) {
  a->critical--;
  g_open_critical.fetch_sub(1);
}
const char* JNIEnv::GetStringUTFChars(jstring s, jboolean*) { return s->utf; }
void JNIEnv::ReleaseStringUTFChars(jstring, const char*) {}
jstring JNIEnv::NewStringUTF(const char* c) {
  auto* s = new _jstring();
  s->utf = strdup(c ? c : "");
  return s;
}
jbyteArray JNIEnv::NewByteArray(jsize n) {
  auto* a = new _jbyteArray();
  a->data = calloc((size_t)(n > 0 ? n : 1), 1);
  a->len = n;
  return a;
}
void JNIEnv::SetByteArrayRegion(jbyteArray a, jsize start, jsize n, const jbyte* buf) {
  memcpy((char*)a->data + start, buf, (size_t)n);
}

// ---- the shim's entry points (declared like javah would)
#define FN(ret, name) extern "C" ret Java_com_fenbi_mp4j_comm_Native_##name
FN(jstring, lastError)(JNIEnv*, jclass);
FN(jlong, masterCreate)(JNIEnv*, jclass, jint, jint, jstring);
FN(jint, masterStart)(JNIEnv*, jclass, jlong);
FN(jint, masterPort)(JNIEnv*, jclass, jlong);
FN(jint, masterWait)(JNIEnv*, jclass, jlong);
FN(void, masterDestroy)(JNIEnv*, jclass, jlong);
FN(jlong, commCreate)(JNIEnv*, jclass, jstring, jstring, jint, jint);
FN(void, commDestroy)(JNIEnv*, jclass, jlong);
FN(jint, commRank)(JNIEnv*, jclass, jlong);
FN(jint, commSize)(JNIEnv*, jclass, jlong);
FN(jint, commBarrier)(JNIEnv*, jclass, jlong);
FN(jint, commClose)(JNIEnv*, jclass, jlong, jint);
FN(jint, hostAllreduce)(JNIEnv*, jclass, jlong, jobject, jint, jint, jlong, jlong);
FN(jint, hostReduceScatter)(JNIEnv*, jclass, jlong, jobject, jint, jint, jlong, jlongArray);
FN(jint, hostAllgather)(JNIEnv*, jclass, jlong, jobject, jint, jlongArray, jlongArray);
FN(jint, hostBroadcast)(JNIEnv*, jclass, jlong, jobject, jint, jlong, jlong, jint);
FN(jbyteArray, sendrecvBytes)(JNIEnv*, jclass, jlong, jint, jbyteArray, jint);
FN(jlong, tgroupCreate)(JNIEnv*, jclass, jint);
FN(void, tgroupDestroy)(JNIEnv*, jclass, jlong);
FN(jint, threadGpuAllreduce)(JNIEnv*, jclass, jlong, jlong, jint, jlong, jint, jint, jlong, jlong, jlong);
#define CALL(name) Java_com_fenbi_mp4j_comm_Native_##name

template <typename T>
struct JArr : _jarray {
  std::vector<T> v;
  explicit JArr(size_t n, T fill = T()) : v(n, fill) {
    data = v.data();
    len = (jsize)n;
  }
  T& operator[](size_t i) { return v[i]; }
};
struct JLongs : _jlongArray {
  std::vector<jlong> v;
  explicit JLongs(std::vector<jlong> x) : v(std::move(x)) {
    data = v.data();
    len = (jsize)v.size();
  }
};
static jstring jstr(const char* c) {
  auto* s = new _jstring();
  s->utf = c;
  return s;
}

static void slave(int p, int port, int hint) {
  JNIEnv env;
  jlong h = CALL(commCreate)(&env, nullptr, jstr("jni"), jstr("127.0.0.1"), port, hint);
  if (!h) {
    CHECK(false, "commCreate: %s", CALL(lastError)(&env, nullptr)->utf);
    return;
  }
  const int r = CALL(commRank)(&env, nullptr, h);
  CHECK(CALL(commSize)(&env, nullptr, h) == p, "size");
  const int n = 100003;
  {  // double[] allreduce SUM over [1, n - 2): the Java array itself must hold the result afterwards
    JArr<double> a((size_t)n, (double)(r + 1));
    CHECK(CALL(hostAllreduce)(&env, nullptr, h, &a, MP4J_F64, MP4J_SUM, 1, n - 2) == 0, "allreduce: %s",
          CALL(lastError)(&env, nullptr)->utf);
    const double want = p * (p + 1) / 2.0;
    CHECK(a[0] == r + 1 && a[1] == want && a[(size_t)n - 3] == want && a[(size_t)n - 2] == r + 1, "allreduce values %f %f", a[1],
          a[(size_t)n - 2]);
    CHECK(a.critical == 0, "critical region left open");
  }
  {  // float[] reduce-scatter MAX + allgather
    JArr<float> a((size_t)n, (float)r);
    std::vector<jlong> cnt((size_t)p, n / p), fr((size_t)p), to((size_t)p);
    for (int q = 0; q < p; ++q) fr[(size_t)q] = (jlong)q * (n / p), to[(size_t)q] = fr[(size_t)q] + n / p;
    JLongs jc(cnt), jf(fr), jt(to);
    CHECK(CALL(hostReduceScatter)(&env, nullptr, h, &a, MP4J_F32, MP4J_MAX, 0, &jc) == 0, "reduce-scatter");
    CHECK(a[(size_t)fr[(size_t)r]] == (float)(p - 1), "reduce-scatter value %f", a[(size_t)fr[(size_t)r]]);
    CHECK(CALL(hostAllgather)(&env, nullptr, h, &a, MP4J_F32, &jf, &jt) == 0, "allgather");
    CHECK(a[0] == (float)(p - 1) && a[(size_t)to[(size_t)p - 1] - 1] == (float)(p - 1), "allgather values");
  }
  {  // long[] broadcast from the last rank
    JArr<int64_t> a(1000, r == p - 1 ? 42 : 0);
    CHECK(CALL(hostBroadcast)(&env, nullptr, h, &a, MP4J_I64, 10, 990, p - 1) == 0, "broadcast");
    CHECK(a[10] == 42 && a[989] == 42 && (r == p - 1 || a[9] == 0), "broadcast values");
  }
  {  // byte transport ring: send own rank to the right, receive the left neighbour's
    _jbyteArray msg;
    char payload[4] = {(char)r, 1, 2, 3};
    msg.data = payload;
    msg.len = 4;
    jbyteArray got = CALL(sendrecvBytes)(&env, nullptr, h, (r + 1) % p, &msg, (r + p - 1) % p);
    CHECK(got && got->len == 4 && ((char*)got->data)[0] == (char)((r + p - 1) % p), "sendrecvBytes");
  }
  {  // an illegal call fails with a message instead of crashing
    JArr<double> a(16, 1.0);
    CHECK(CALL(hostAllreduce)(&env, nullptr, h, &a, MP4J_F64, MP4J_SUM, 8, 4) != 0, "from > to must fail");
    CHECK(strlen(CALL(lastError)(&env, nullptr)->utf) > 0, "lastError empty");
  }
  CHECK(CALL(commBarrier)(&env, nullptr, h) == 0, "barrier");
  CHECK(CALL(commClose)(&env, nullptr, h, 0) == 0, "close");
  CALL(commDestroy)(&env, nullptr, h);
}

int main(int argc, char** argv) {
  const int p = argc > 1 ? atoi(argv[1]) : 3;
  JNIEnv env;
  jlong m = CALL(masterCreate)(&env, nullptr, p, 0, jstr(""));
  if (!m || CALL(masterStart)(&env, nullptr, m) != 0) {
    fprintf(stderr, "master: %s\n", CALL(lastError)(&env, nullptr)->utf);
    return 2;
  }
  const int port = CALL(masterPort)(&env, nullptr, m);
  std::vector<std::thread> th;
  for (int i = 0; i < p; ++i) th.emplace_back(slave, p, port, i);
  for (auto& t : th) t.join();
  const int code = CALL(masterWait)(&env, nullptr, m);
  CHECK(code == 0, "master exit code %d", code);
  CALL(masterDestroy)(&env, nullptr, m);
  {  // thread group handle of the thread-level device methods: created and destroyed without a GPU; a call with a
     // thread id outside the group fails with a message instead of touching the (absent) GPU plane
    jlong tg = CALL(tgroupCreate)(&env, nullptr, 4);
    CHECK(tg != 0, "tgroupCreate");
    CHECK(CALL(threadGpuAllreduce)(&env, nullptr, tg, 0, 7, 0, MP4J_F32, MP4J_SUM, 0, 16, 0) != 0, "bad thread id must fail");
    CHECK(strstr(CALL(lastError)(&env, nullptr)->utf, "threadId") != nullptr, "error text: %s", CALL(lastError)(&env, nullptr)->utf);
    CALL(tgroupDestroy)(&env, nullptr, tg);
  }
  CHECK(g_open_critical.load() == 0, "critical regions left open");
  if (g_fail.load() == 0) printf("jni harness ok (p=%d)\n", p);
  return g_fail.load() ? 1 : 0;
}
