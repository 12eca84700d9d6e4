package com.fenbi.mp4j.comm;

/** JNI bridge to libmp4j.so (jni/mp4j_jni.cc). Every method maps 1:1 onto a function of
 *  include/mp4j.h; a negative return is turned into Mp4jException(lastError()) by the callers. */
final class Native {
    static { System.loadLibrary("mp4j_jni"); }
    private Native() {}

    static native String lastError();
    // master
    static native long masterCreate(int slaveNum, int port, String logDir);
    static native int masterStart(long h);
    static native int masterPort(long h);
    static native int masterWait(long h);      // blocks; returns the job exit code
    static native void masterDestroy(long h);
    // communicator
    static native long commCreate(String login, String masterHost, int masterPort, int rankHint);
    static native void commDestroy(long h);
    static native int commRank(long h);
    static native int commSize(long h);
    static native int commBarrier(long h);
    static native int commLog(long h, int level, String msg);
    static native int commClose(long h, int code);
    static native int commWriteFile(long h, String content, String fileName);
    // host collectives on primitive arrays: `arr` is double[]/float[]/long[]/int[]/short[]/byte[]
    static native int hostReduceScatter(long h, Object arr, int dtype, int op, long from, long[] counts);
    static native int hostAllgather(long h, Object arr, int dtype, long[] froms, long[] tos);
    static native int hostGather(long h, Object arr, int dtype, long[] froms, long[] tos, int root);
    static native int hostScatter(long h, Object arr, int dtype, long[] froms, long[] tos, int root);
    static native int hostBroadcast(long h, Object arr, int dtype, long from, long to, int root);
    static native int hostReduce(long h, Object arr, int dtype, int op, long from, long to, int root);
    static native int hostAllreduce(long h, Object arr, int dtype, int op, long from, long to);
    static native int hostAllreduceRpc(long h, Object arr, int dtype, int op, long n);
    // byte transport for String/Object/map payloads
    static native byte[] sendrecvBytes(long h, int dest, byte[] data, int src);
    static native int sendBytes(long h, int dest, byte[] data);
    static native byte[] recvBytes(long h, int src);
    // GPU plane on raw device addresses (e.g. from a CUDA binding of the caller's choice)
    static native long gpuInit(long h, int device, long heapBytes, long stagingBytes);
    static native int gpuAllreduce(long g, long devPtr, int dtype, int op, long from, long to, int algo, long stream);
    static native int gpuReduceScatter(long g, long devPtr, int dtype, int op, long from, long[] counts, long stream);
    static native int gpuAllgather(long g, long devPtr, int dtype, long[] froms, long[] tos, long stream);
    static native int gpuBroadcast(long g, long devPtr, int dtype, long from, long to, int root, long stream);
    // ThreadCommSlave on device buffers: every thread calls with its own buffer, the rendezvous is native
    static native long tgroupCreate(int threadNum);
    static native void tgroupDestroy(long tg);
    static native int threadGpuAllreduce(long tg, long g, int tid, long devPtr, int dtype, int op, long from, long to, long stream);
    static native int threadGpuSegments(long tg, long g, int tid, int launchTid, long devPtr, int dtype, int op, long[] segs, String name, long stream);
}
