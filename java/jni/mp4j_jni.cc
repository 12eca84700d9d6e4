// JNI shim over the C ABI of include/mp4j.h for the com.fenbi.mp4j Java facade.
// Not built on this image (no JDK / jni.h): see java/build.sh. The same C ABI is exercised from
// Python (ctypes) by the test-suite, so the surface below is covered even where Java is absent.
#include <jni.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/mp4j.h"

namespace {
struct Str {
  JNIEnv* e;
  jstring s;
  const char* c;
  Str(JNIEnv* env, jstring js) : e(env), s(js), c(js ? env->GetStringUTFChars(js, nullptr) : nullptr) {}
  ~Str() {
    if (c) e->ReleaseStringUTFChars(s, c);
  }
};
struct Longs {
  JNIEnv* e;
  jlongArray a;
  jlong* p;
  Longs(JNIEnv* env, jlongArray arr) : e(env), a(arr), p(env->GetLongArrayElements(arr, nullptr)) {}
  ~Longs() { e->ReleaseLongArrayElements(a, p, JNI_ABORT); }
  const int64_t* get() const { return (const int64_t*)p; }
};
// A native copy of a primitive Java array for the duration of a host collective. The collective blocks on
// sockets, futexes and master RPCs, and JNI forbids blocking inside a Get/ReleasePrimitiveArrayCritical
// region (any other Java thread that triggers a GC would stall behind it - with T threads of a
// ThreadCommSlave that can deadlock the job). So the critical pin is taken only around the two memcpys:
// copy in, run the collective on the native buffer, copy the result back.
static size_t elem_size(int dt) {
  switch (dt) {
    case MP4J_F64: case MP4J_I64: return 8;
    case MP4J_F32: case MP4J_I32: return 4;
    case MP4J_I16: case MP4J_BF16: case MP4J_F16: return 2;
    default: return 1;
  }
}
struct Pinned {
  JNIEnv* e;
  jarray a;
  void* p;
  size_t bytes;
  Pinned(JNIEnv* env, jobject arr, int dt) : e(env), a((jarray)arr), p(nullptr), bytes(0) {
    bytes = (size_t)env->GetArrayLength(a) * elem_size(dt);
    p = malloc(bytes ? bytes : 1);
    void* j = env->GetPrimitiveArrayCritical(a, nullptr);
    if (j) {
      memcpy(p, j, bytes);
      env->ReleasePrimitiveArrayCritical(a, j, JNI_ABORT);  // nothing written to the Java array yet
    }
  }
  ~Pinned() {
    void* j = e->GetPrimitiveArrayCritical(a, nullptr);
    if (j) {
      memcpy(j, p, bytes);
      e->ReleasePrimitiveArrayCritical(a, j, 0);
    }
    free(p);
  }
};
jbyteArray to_bytes(JNIEnv* env, char* data, size_t n) {
  jbyteArray out = env->NewByteArray((jsize)n);
  env->SetByteArrayRegion(out, 0, (jsize)n, (const jbyte*)data);
  mp4j_free(data);
  return out;
}
}  // namespace

#define FN(ret, name) extern "C" JNIEXPORT ret JNICALL Java_com_fenbi_mp4j_comm_Native_##name

FN(jstring, lastError)(JNIEnv* env, jclass) { return env->NewStringUTF(mp4j_last_error()); }

FN(jlong, masterCreate)(JNIEnv* env, jclass, jint n, jint port, jstring logDir) {
  Str d(env, logDir);
  mp4j_master* m = nullptr;
  if (mp4j_master_create(n, port, 0, 0, 0, 0, d.c, nullptr, 0, &m) != 0) return 0;
  return (jlong)m;
}
FN(jint, masterStart)(JNIEnv*, jclass, jlong h) { return mp4j_master_start((mp4j_master*)h); }
FN(jint, masterPort)(JNIEnv*, jclass, jlong h) { return mp4j_master_port((mp4j_master*)h); }
FN(jint, masterWait)(JNIEnv*, jclass, jlong h) {
  int code = 1;
  mp4j_master_wait((mp4j_master*)h, -1, &code);
  return code;
}
FN(void, masterDestroy)(JNIEnv*, jclass, jlong h) { mp4j_master_destroy((mp4j_master*)h); }

FN(jlong, commCreate)(JNIEnv* env, jclass, jstring login, jstring host, jint port, jint hint) {
  Str l(env, login), h(env, host);
  mp4j_comm* c = nullptr;
  if (mp4j_comm_create(l.c, h.c, port, hint, nullptr, &c) != 0) return 0;
  return (jlong)c;
}
FN(void, commDestroy)(JNIEnv*, jclass, jlong h) { mp4j_comm_destroy((mp4j_comm*)h); }
FN(jint, commRank)(JNIEnv*, jclass, jlong h) { return mp4j_comm_rank((mp4j_comm*)h); }
FN(jint, commSize)(JNIEnv*, jclass, jlong h) { return mp4j_comm_size((mp4j_comm*)h); }
FN(jint, commBarrier)(JNIEnv*, jclass, jlong h) { return mp4j_comm_barrier((mp4j_comm*)h); }
FN(jint, commLog)(JNIEnv* env, jclass, jlong h, jint lvl, jstring msg) {
  Str m(env, msg);
  return mp4j_comm_log((mp4j_comm*)h, lvl, m.c);
}
FN(jint, commClose)(JNIEnv*, jclass, jlong h, jint code) { return mp4j_comm_close((mp4j_comm*)h, code); }
FN(jint, commWriteFile)(JNIEnv* env, jclass, jlong h, jstring content, jstring name) {
  Str c(env, content), n(env, name);
  return mp4j_comm_write_file((mp4j_comm*)h, c.c, n.c);
}

FN(jint, hostReduceScatter)(JNIEnv* env, jclass, jlong h, jobject arr, jint dt, jint op, jlong from, jlongArray counts) {
  Longs c(env, counts);
  Pinned p(env, arr, dt);
  return mp4j_host_reduce_scatter((mp4j_comm*)h, p.p, dt, op, from, c.get());
}
FN(jint, hostAllgather)(JNIEnv* env, jclass, jlong h, jobject arr, jint dt, jlongArray froms, jlongArray tos) {
  Longs f(env, froms), t(env, tos);
  Pinned p(env, arr, dt);
  return mp4j_host_allgather((mp4j_comm*)h, p.p, dt, f.get(), t.get());
}
FN(jint, hostGather)(JNIEnv* env, jclass, jlong h, jobject arr, jint dt, jlongArray froms, jlongArray tos, jint root) {
  Longs f(env, froms), t(env, tos);
  Pinned p(env, arr, dt);
  return mp4j_host_gather((mp4j_comm*)h, p.p, dt, f.get(), t.get(), root);
}
FN(jint, hostScatter)(JNIEnv* env, jclass, jlong h, jobject arr, jint dt, jlongArray froms, jlongArray tos, jint root) {
  Longs f(env, froms), t(env, tos);
  Pinned p(env, arr, dt);
  return mp4j_host_scatter((mp4j_comm*)h, p.p, dt, f.get(), t.get(), root);
}
FN(jint, hostBroadcast)(JNIEnv* env, jclass, jlong h, jobject arr, jint dt, jlong from, jlong to, jint root) {
  Pinned p(env, arr, dt);
  return mp4j_host_broadcast((mp4j_comm*)h, p.p, dt, from, to, root);
}
FN(jint, hostReduce)(JNIEnv* env, jclass, jlong h, jobject arr, jint dt, jint op, jlong from, jlong to, jint root) {
  Pinned p(env, arr, dt);
  return mp4j_host_reduce((mp4j_comm*)h, p.p, dt, op, from, to, root);
}
FN(jint, hostAllreduce)(JNIEnv* env, jclass, jlong h, jobject arr, jint dt, jint op, jlong from, jlong to) {
  Pinned p(env, arr, dt);
  return mp4j_host_allreduce((mp4j_comm*)h, p.p, dt, op, from, to);
}
FN(jint, hostAllreduceRpc)(JNIEnv* env, jclass, jlong h, jobject arr, jint dt, jint op, jlong n) {
  Pinned p(env, arr, dt);
  return mp4j_host_allreduce_rpc((mp4j_comm*)h, p.p, dt, op, n);
}

FN(jbyteArray, sendrecvBytes)(JNIEnv* env, jclass, jlong h, jint dest, jbyteArray data, jint src) {
  jsize n = env->GetArrayLength(data);
  jbyte* d = env->GetByteArrayElements(data, nullptr);
  char* out = nullptr;
  size_t out_len = 0;
  int rc = mp4j_comm_sendrecv_bytes((mp4j_comm*)h, dest, d, (size_t)n, src, &out, &out_len);
  env->ReleaseByteArrayElements(data, d, JNI_ABORT);
  return rc == 0 ? to_bytes(env, out, out_len) : nullptr;
}
FN(jint, sendBytes)(JNIEnv* env, jclass, jlong h, jint dest, jbyteArray data) {
  jsize n = env->GetArrayLength(data);
  jbyte* d = env->GetByteArrayElements(data, nullptr);
  int rc = mp4j_comm_send_bytes((mp4j_comm*)h, dest, d, (size_t)n);
  env->ReleaseByteArrayElements(data, d, JNI_ABORT);
  return rc;
}
FN(jbyteArray, recvBytes)(JNIEnv* env, jclass, jlong h, jint src) {
  char* out = nullptr;
  size_t out_len = 0;
  return mp4j_comm_recv_bytes((mp4j_comm*)h, src, &out, &out_len) == 0 ? to_bytes(env, out, out_len) : nullptr;
}

FN(jlong, gpuInit)(JNIEnv*, jclass, jlong h, jint device, jlong heap, jlong staging) {
  mp4j_gpu* g = nullptr;
  if (mp4j_gpu_init((mp4j_comm*)h, device, (size_t)heap, (size_t)staging, &g) != 0) return 0;
  return (jlong)g;
}
FN(jint, gpuAllreduce)(JNIEnv*, jclass, jlong g, jlong ptr, jint dt, jint op, jlong from, jlong to, jint algo, jlong stream) {
  void* arrs[1] = {(void*)ptr};
  return mp4j_gpu_allreduce((mp4j_gpu*)g, arrs, dt, op, from, to, algo, (void*)stream);
}
FN(jint, gpuReduceScatter)(JNIEnv* env, jclass, jlong g, jlong ptr, jint dt, jint op, jlong from, jlongArray counts, jlong stream) {
  Longs c(env, counts);
  void* arrs[1] = {(void*)ptr};
  return mp4j_gpu_reduce_scatter((mp4j_gpu*)g, arrs, dt, op, from, c.get(), 0, (void*)stream);
}
FN(jint, gpuAllgather)(JNIEnv* env, jclass, jlong g, jlong ptr, jint dt, jlongArray froms, jlongArray tos, jlong stream) {
  Longs f(env, froms), t(env, tos);
  void* arrs[1] = {(void*)ptr};
  return mp4j_gpu_allgather((mp4j_gpu*)g, arrs, dt, f.get(), t.get(), 0, (void*)stream);
}
FN(jint, gpuBroadcast)(JNIEnv*, jclass, jlong g, jlong ptr, jint dt, jlong from, jlong to, jint root, jlong stream) {
  void* arrs[1] = {(void*)ptr};
  return mp4j_gpu_broadcast((mp4j_gpu*)g, arrs, dt, from, to, root, 0, (void*)stream);
}

// ---- ThreadCommSlave on device buffers: every Java thread calls with its own buffer (include/mp4j_gpu.h)
FN(jlong, tgroupCreate)(JNIEnv*, jclass, jint threadNum) {
  mp4j_tgroup* tg = nullptr;
  if (mp4j_tgroup_create(threadNum, &tg) != 0) return 0;
  return (jlong)tg;
}
FN(void, tgroupDestroy)(JNIEnv*, jclass, jlong tg) { mp4j_tgroup_destroy((mp4j_tgroup*)tg); }
FN(jint, threadGpuAllreduce)(JNIEnv*, jclass, jlong tg, jlong g, jint tid, jlong ptr, jint dt, jint op, jlong from, jlong to, jlong stream) {
  return mp4j_thread_gpu_allreduce((mp4j_tgroup*)tg, (mp4j_gpu*)g, tid, (void*)ptr, dt, op, from, to, 1, 0, 0, 0, 0, (void*)stream);
}
FN(jint, threadGpuSegments)(JNIEnv* env, jclass, jlong tg, jlong g, jint tid, jint launchTid, jlong ptr, jint dt, jint op, jlongArray segs, jstring name, jlong stream) {
  Longs sg(env, segs);
  Str nm(env, name);
  return mp4j_thread_gpu_segments((mp4j_tgroup*)tg, (mp4j_gpu*)g, tid, launchTid, (void*)ptr, dt, op, sg.get(),
                                  (int)(env->GetArrayLength(segs) / 6), nm.c, (void*)stream);
}
