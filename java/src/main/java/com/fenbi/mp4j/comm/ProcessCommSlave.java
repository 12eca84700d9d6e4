package com.fenbi.mp4j.comm;

import com.fenbi.mp4j.exception.Mp4jException;
import com.fenbi.mp4j.operand.Operand;
import com.fenbi.mp4j.operator.IObjectOperator;
import com.fenbi.mp4j.operator.IOperator;

import java.io.ByteArrayInputStream;
import java.io.ByteArrayOutputStream;
import java.io.ObjectInputStream;
import java.io.ObjectOutputStream;
import java.io.PrintWriter;
import java.io.StringWriter;
import java.util.ArrayList;
import java.util.HashMap;
import java.util.List;
import java.util.Map;

/** Process-level communicator with the reference's method names. Primitive arrays go to the native
 *  host collectives (or, for device addresses, the GPU plane); maps and objects use the same ring
 *  algorithms written here over the native byte transport with Java serialization as the codec. */
public class ProcessCommSlave {
    private final long handle;
    private final int rank;
    private final int slaveNum;
    private long gpu = 0;

    public ProcessCommSlave(String loginName, String masterHost, int masterPort) throws Mp4jException {
        handle = Native.commCreate(loginName, masterHost, masterPort, -1);
        if (handle == 0) throw new Mp4jException(Native.lastError());
        rank = Native.commRank(handle);
        slaveNum = Native.commSize(handle);
    }

    public int getRank() { return rank; }
    public int getSlaveNum() { return slaveNum; }
    private static void check(int rc) throws Mp4jException { if (rc != 0) throw new Mp4jException(Native.lastError()); }

    public void barrier() throws Mp4jException { check(Native.commBarrier(handle)); }
    public void info(String info) throws Mp4jException { info(info, true); }
    public void info(String info, boolean onlyRank0) throws Mp4jException {
        if (onlyRank0 && rank != 0) return;
        check(Native.commLog(handle, 1, "[rank=" + rank + "] " + info));
    }
    public void debug(String debug) throws Mp4jException { debug(debug, true); }
    public void debug(String debug, boolean onlyRank0) throws Mp4jException {
        if (onlyRank0 && rank != 0) return;
        check(Native.commLog(handle, 0, "[rank=" + rank + "] " + debug));
    }
    public void error(String error) throws Mp4jException { check(Native.commLog(handle, 2, "[rank=" + rank + "] " + error)); }
    public void exception(Exception e) throws Mp4jException {
        StringWriter sw = new StringWriter();
        e.printStackTrace(new PrintWriter(sw));
        try { error("slave exception:" + sw); } finally {
            try { Thread.sleep(5000); } catch (InterruptedException ignored) { }
            close(1);
        }
    }
    public void writeFile(String content, String fileName) throws Mp4jException { check(Native.commWriteFile(handle, content, fileName)); }
    public void close(int code) throws Mp4jException { check(Native.commClose(handle, code)); }

    private static long[] toLong(int[] a) { long[] r = new long[a.length]; for (int i = 0; i < a.length; i++) r[i] = a[i]; return r; }

    // ---- primitive arrays (T = double[] | float[] | long[] | int[] | short[] | byte[])
    public <T> T gatherArray(T arr, Operand operand, int[] froms, int[] tos, int rootRank) throws Mp4jException {
        if (slaveNum > 1) check(Native.hostGather(handle, arr, operand.dtype, toLong(froms), toLong(tos), rootRank));
        return arr;
    }
    public <T> T allgatherArray(T arr, Operand operand, int[] froms, int[] tos) throws Mp4jException {
        if (slaveNum > 1) check(Native.hostAllgather(handle, arr, operand.dtype, toLong(froms), toLong(tos)));
        return arr;
    }
    public <T> T scatterArray(T arr, Operand operand, int[] froms, int[] tos, int rootRank) throws Mp4jException {
        if (slaveNum > 1) check(Native.hostScatter(handle, arr, operand.dtype, toLong(froms), toLong(tos), rootRank));
        return arr;
    }
    public <T> T broadcastArray(T arr, Operand operand, int from, int to, int rootRank) throws Mp4jException {
        if (slaveNum > 1) check(Native.hostBroadcast(handle, arr, operand.dtype, from, to, rootRank));
        return arr;
    }
    public <T> T reduceScatterArray(T arr, Operand operand, IOperator operator, int from, int[] counts) throws Mp4jException {
        if (slaveNum > 1) check(Native.hostReduceScatter(handle, arr, operand.dtype, nativeOp(operator), from, toLong(counts)));
        return arr;
    }
    public <T> T reduceArray(T arr, Operand operand, IOperator operator, int from, int to, int rootRank) throws Mp4jException {
        if (slaveNum > 1) check(Native.hostReduce(handle, arr, operand.dtype, nativeOp(operator), from, to, rootRank));
        return arr;
    }
    public <T> T allreduceArray(T arr, Operand operand, IOperator operator, int from, int to) throws Mp4jException {
        if (slaveNum > 1) check(Native.hostAllreduce(handle, arr, operand.dtype, nativeOp(operator), from, to));
        return arr;
    }
    public <T> T allreduceArrayRpc(T arr, Operand operand, IOperator operator) throws Mp4jException {
        if (slaveNum > 1) check(Native.hostAllreduceRpc(handle, arr, operand.dtype, nativeOp(operator), java.lang.reflect.Array.getLength(arr)));
        return arr;
    }
    private static int nativeOp(IOperator op) throws Mp4jException {
        if (op.nativeCode() < 0) throw new Mp4jException("custom operators are supported on maps/objects only in the Java facade");
        return op.nativeCode();
    }

    // ---- device arrays (raw CUDA addresses): the sm_100a kernels
    public void initGpu(int device, long heapBytes, long stagingBytes) throws Mp4jException {
        gpu = Native.gpuInit(handle, device, heapBytes, stagingBytes);
        if (gpu == 0) throw new Mp4jException(Native.lastError());
    }
    public void allreduceDevice(long devPtr, Operand operand, IOperator operator, long from, long to, long stream) throws Mp4jException {
        check(Native.gpuAllreduce(gpu, devPtr, operand.dtype, nativeOp(operator), from, to, 0, stream));
    }
    public void reduceScatterDevice(long devPtr, Operand operand, IOperator operator, long from, long[] counts, long stream) throws Mp4jException {
        check(Native.gpuReduceScatter(gpu, devPtr, operand.dtype, nativeOp(operator), from, counts, stream));
    }
    public void allgatherDevice(long devPtr, Operand operand, long[] froms, long[] tos, long stream) throws Mp4jException {
        check(Native.gpuAllgather(gpu, devPtr, operand.dtype, froms, tos, stream));
    }
    public void broadcastDevice(long devPtr, Operand operand, long from, long to, int root, long stream) throws Mp4jException {
        check(Native.gpuBroadcast(gpu, devPtr, operand.dtype, from, to, root, stream));
    }

    // ---- maps (Map<String,T> with Serializable T) over the byte transport
    private static byte[] ser(Object o) throws Mp4jException {
        try {
            ByteArrayOutputStream bos = new ByteArrayOutputStream();
            ObjectOutputStream oos = new ObjectOutputStream(bos);
            oos.writeObject(o);
            oos.close();
            return bos.toByteArray();
        } catch (Exception e) { throw new Mp4jException(e); }
    }
    @SuppressWarnings("unchecked")
    private static <X> X deser(byte[] b) throws Mp4jException {
        try { return (X) new ObjectInputStream(new ByteArrayInputStream(b)).readObject(); }
        catch (Exception e) { throw new Mp4jException(e); }
    }
    private static int partition(String key, int p) { int m = key.hashCode() % p; return m < 0 ? m + p : m; }

    /** reduceScatterMap: maps.get(i) is this rank's contribution to rank i (ring, p-1 steps). */
    public <T> Map<String, T> reduceScatterMap(List<Map<String, T>> maps, Operand operand, IObjectOperator<T> operator) throws Mp4jException {
        if (slaveNum == 1) return maps.get(0);
        int next = (rank + 1) % slaveNum, prev = (rank - 1 + slaveNum) % slaveNum;
        for (int s = 0; s < slaveNum - 1; s++) {
            int sseg = ((rank - 1 - s) % slaveNum + slaveNum) % slaveNum;
            int rseg = ((rank - 2 - s) % slaveNum + slaveNum) % slaveNum;
            Map<String, T> incoming = deser(Native.sendrecvBytes(handle, next, ser(new HashMap<>(maps.get(sseg))), prev));
            Map<String, T> local = maps.get(rseg);
            for (Map.Entry<String, T> e : incoming.entrySet()) {
                T cur = local.get(e.getKey());
                local.put(e.getKey(), cur == null ? e.getValue() : operator.apply(cur, e.getValue()));
            }
        }
        return maps.get(rank);
    }
    public <T> List<Map<String, T>> allgatherMap(Map<String, T> map, Operand operand) throws Mp4jException {
        List<Map<String, T>> out = new ArrayList<>();
        for (int i = 0; i < slaveNum; i++) out.add(null);
        out.set(rank, map);
        int next = (rank + 1) % slaveNum, prev = (rank - 1 + slaveNum) % slaveNum;
        for (int s = 0; s < slaveNum - 1; s++) {
            int sseg = ((rank - s) % slaveNum + slaveNum) % slaveNum;
            int rseg = ((rank - 1 - s) % slaveNum + slaveNum) % slaveNum;
            out.set(rseg, deser(Native.sendrecvBytes(handle, next, ser(new HashMap<>(out.get(sseg))), prev)));
        }
        return out;
    }
    public <T> Map<String, T> allreduceMap(Map<String, T> map, Operand operand, IObjectOperator<T> operator) throws Mp4jException {
        if (slaveNum == 1) return map;
        List<Map<String, T>> parts = new ArrayList<>();
        for (int i = 0; i < slaveNum; i++) parts.add(new HashMap<String, T>());
        for (Map.Entry<String, T> e : map.entrySet()) parts.get(partition(e.getKey(), slaveNum)).put(e.getKey(), e.getValue());
        Map<String, T> mine = reduceScatterMap(parts, operand, operator);
        Map<String, T> out = new HashMap<>();
        for (Map<String, T> m : allgatherMap(mine, operand)) out.putAll(m);
        return out;
    }
}
