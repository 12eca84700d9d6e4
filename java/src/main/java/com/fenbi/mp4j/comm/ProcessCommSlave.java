package com.fenbi.mp4j.comm;

import com.fenbi.mp4j.exception.Mp4jException;
import com.fenbi.mp4j.operand.ObjectOperand;
import com.fenbi.mp4j.operand.Operand;
import com.fenbi.mp4j.operator.IObjectOperator;
import com.fenbi.mp4j.operator.IOperator;
import com.fenbi.mp4j.utils.ArrayOps;
import com.fenbi.mp4j.utils.CommUtils;

import java.io.PrintWriter;
import java.io.StringWriter;
import java.util.ArrayList;
import java.util.HashMap;
import java.util.HashSet;
import java.util.List;
import java.util.Map;
import java.util.Set;

/** Process-level communicator with the reference's method names and argument order
 *  (J/comm/ProcessCommSlave.java:143-2211).
 *
 *  Three data paths behind one API, as in the Python binding:
 *   - primitive arrays with a predefined operator: the native host collectives of libmp4j (shared-memory
 *     windows on one machine, TCP ring / tree otherwise), in place on the Java array;
 *   - device addresses: the sm_100a kernels of the GPU plane (the ...Device methods);
 *   - everything else (String / Object arrays, maps, user-implemented operators): the reference's four
 *     algorithms - ring reduce-scatter, ring allgather, binomial gather, binomial scatter - written here
 *     over the native byte transport, payloads serialized by the operand.
 *  A communicator runs one collective at a time (SURVEY 2.7-12); methods are synchronized. */
public class ProcessCommSlave {
    private final long handle;
    private final int rank;
    private final int slaveNum;
    private long gpu = 0;
    private volatile boolean closed = false;

    public ProcessCommSlave(String loginName, String masterHost, int masterPort) throws Mp4jException {
        handle = Native.commCreate(loginName, masterHost, masterPort, -1);
        if (handle == 0) throw new Mp4jException(Native.lastError());
        rank = Native.commRank(handle);
        slaveNum = Native.commSize(handle);
    }

    public int getRank() { return rank; }
    public int getSlaveNum() { return slaveNum; }
    public boolean isClosed() { return closed; }
    private static void check(int rc) throws Mp4jException { if (rc != 0) throw new Mp4jException(Native.lastError()); }

    // ------------------------------------------------------------------ sync / logging / lifecycle
    public void barrier() throws Mp4jException { if (slaveNum > 1) check(Native.commBarrier(handle)); }
    /** default: only rank 0 forwards the message to the master (PCS:308-310) */
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
    /** forwards the stack trace, waits 5 s for the log to drain, closes with code 1 (PCS:360) */
    public void exception(Exception e) throws Mp4jException {
        StringWriter sw = new StringWriter();
        e.printStackTrace(new PrintWriter(sw));
        try {
            error("slave exception:" + sw);
        } finally {
            try { Thread.sleep(5000); } catch (InterruptedException ignored) { }
            close(1);
        }
    }
    /** the file is created on the MASTER (PCS:279) */
    public void writeFile(String content, String fileName) throws Mp4jException { check(Native.commWriteFile(handle, content, fileName)); }
    public synchronized void close(int code) throws Mp4jException {
        if (closed) return;
        closed = true;
        check(Native.commClose(handle, code));
    }

    // ------------------------------------------------------------------ generic transport
    private Object sendrecv(int dest, Object payload, int src, Operand operand) throws Mp4jException {
        byte[] in = Native.sendrecvBytes(handle, dest, operand.toBytes(payload), src);
        if (in == null) throw new Mp4jException(Native.lastError());
        return operand.fromBytes(in);
    }
    private void send(int dest, Object payload, Operand operand) throws Mp4jException {
        check(Native.sendBytes(handle, dest, operand.toBytes(payload)));
    }
    private Object recv(int src, Operand operand) throws Mp4jException {
        byte[] in = Native.recvBytes(handle, src);
        if (in == null) throw new Mp4jException(Native.lastError());
        return operand.fromBytes(in);
    }

    /** op(local, incoming) on whole payloads (array slices or maps) */
    private interface Merger { Object merge(Object local, Object incoming) throws Mp4jException; }

    private static Merger arrayMerger(final IOperator op) {
        return new Merger() {
            @Override public Object merge(Object local, Object incoming) throws Mp4jException {
                ArrayOps.reduce(local, incoming, 0, ArrayOps.length(local), 0, op);
                return local;
            }
        };
    }
    @SuppressWarnings("unchecked")
    private static Merger mapMerger(final IOperator op) {
        return new Merger() {
            @Override public Object merge(Object local, Object incoming) throws Mp4jException {
                ArrayOps.mergeMap((Map<String, Object>) local, (Map<String, Object>) incoming, op);
                return local;
            }
        };
    }

    private int mod(int x) { return ((x % slaveNum) + slaveNum) % slaveNum; }

    /** parts[i] = this rank's contribution to rank i's result; p-1 ring steps (PCS:1329-1373) */
    private Object ringReduceScatter(Object[] parts, Merger merge, Operand operand) throws Mp4jException {
        int next = mod(rank + 1), prev = mod(rank - 1);
        for (int s = 0; s < slaveNum - 1; s++) {
            int sseg = mod(rank - 1 - s), rseg = mod(rank - 2 - s);
            Object incoming = sendrecv(next, parts[sseg], prev, operand);
            parts[rseg] = merge.merge(parts[rseg], incoming);
        }
        return parts[rank];
    }

    /** returns the p payloads indexed by origin rank; p-1 ring steps (PCS:694-737) */
    private Object[] ringAllgather(Object mine, Operand operand) throws Mp4jException {
        int next = mod(rank + 1), prev = mod(rank - 1);
        Object[] out = new Object[slaveNum];
        out[rank] = mine;
        for (int s = 0; s < slaveNum - 1; s++) {
            int sseg = mod(rank - s), rseg = mod(rank - 1 - s);
            out[rseg] = sendrecv(next, out[sseg], prev, operand);
        }
        return out;
    }

    /** binomial gather towards root on relative ranks; the returned map is complete on the root */
    @SuppressWarnings("unchecked")
    private Map<Integer, Object> treeGather(Object mine, int root, Operand operand) throws Mp4jException {
        int vr = mod(rank - root);
        Map<Integer, Object> acc = new HashMap<Integer, Object>();
        acc.put(rank, mine);
        for (int m = 1; m < slaveNum; m <<= 1) {
            if ((vr & m) != 0) {
                send(mod(vr - m + root), acc, operand);
                break;
            }
            if (vr + m < slaveNum) acc.putAll((Map<Integer, Object>) recv(mod(vr + m + root), operand));
        }
        return acc;
    }

    /** binomial scatter of payloads[0..p) held by root; returns this rank's payload */
    @SuppressWarnings("unchecked")
    private Object treeScatter(Object[] payloads, int root, Operand operand) throws Mp4jException {
        int vr = mod(rank - root);
        int top = 1;
        while (top < slaveNum) top <<= 1;
        top >>= 1;
        Map<Integer, Object> acc;
        int m;
        if (vr == 0) {
            acc = new HashMap<Integer, Object>();
            for (int q = 0; q < slaveNum; q++) acc.put(q, payloads[q]);
            m = top;
        } else {
            int lsb = vr & -vr;
            acc = (Map<Integer, Object>) recv(mod(vr - lsb + root), operand);
            m = lsb >> 1;
        }
        for (; m >= 1; m >>= 1) {
            if (vr + m >= slaveNum) continue;
            Map<Integer, Object> sub = new HashMap<Integer, Object>();
            for (int v = vr + m; v < Math.min(vr + 2 * m, slaveNum); v++) sub.put(mod(v + root), acc.get(mod(v + root)));
            send(mod(vr + m + root), sub, operand);
        }
        return acc.get(rank);
    }

    private void checkRoot(int root) throws Mp4jException {
        if (root < 0 || root >= slaveNum) throw new Mp4jException("rootRank " + root + " out of range [0," + slaveNum + ")");
    }
    private static long[] toLong(int[] a) { long[] r = new long[a.length]; for (int i = 0; i < a.length; i++) r[i] = a[i]; return r; }
    private static boolean nativePath(Operand operand, IOperator op) {
        return operand.isPrimitive() && (op == null || op.nativeCode() >= 0);
    }
    private static boolean isNativeArray(Object arr) {
        return arr instanceof double[] || arr instanceof float[] || arr instanceof long[] || arr instanceof int[]
                || arr instanceof short[] || arr instanceof byte[];
    }

    // ------------------------------------------------------------------ arrays
    // T = double[] | float[] | long[] | int[] | short[] | byte[] | String[] | E[]; all in place.

    /** rank i's [sendfroms[i], sendtos[i]) lands at the same offsets of root's array (PCS:535) */
    public synchronized <T> T gatherArray(T arrData, Operand operand, int[] sendfroms, int[] sendtos, int rootRank) throws Mp4jException {
        CommUtils.isfromsTosLegal(sendfroms, sendtos);
        CommUtils.isRangeInArray(arrData, CommUtils.maxEnd(sendtos));
        checkRoot(rootRank);
        if (slaveNum == 1) return arrData;
        if (nativePath(operand, null) && isNativeArray(arrData)) {
            check(Native.hostGather(handle, arrData, operand.dtype, toLong(sendfroms), toLong(sendtos), rootRank));
            return arrData;
        }
        Map<Integer, Object> acc = treeGather(ArrayOps.slice(arrData, sendfroms[rank], sendtos[rank]), rootRank, operand);
        if (rank == rootRank)
            for (Map.Entry<Integer, Object> e : acc.entrySet())
                if (e.getKey() != rank) ArrayOps.put(arrData, sendfroms[e.getKey()], e.getValue());
        return arrData;
    }

    /** every rank ends with all p segments (PCS:620) */
    public synchronized <T> T allgatherArray(T arrData, Operand operand, int[] froms, int[] tos) throws Mp4jException {
        CommUtils.isfromsTosLegal(froms, tos);
        CommUtils.isRangeInArray(arrData, CommUtils.maxEnd(tos));
        if (slaveNum == 1) return arrData;
        if (nativePath(operand, null) && isNativeArray(arrData)) {
            check(Native.hostAllgather(handle, arrData, operand.dtype, toLong(froms), toLong(tos)));
            return arrData;
        }
        Object[] all = ringAllgather(ArrayOps.slice(arrData, froms[rank], tos[rank]), operand);
        for (int q = 0; q < slaveNum; q++) if (q != rank) ArrayOps.put(arrData, froms[q], all[q]);
        return arrData;
    }

    /** root's [from, to) to everybody = scatter of p balanced chunks + allgather (PCS:750-770) */
    public synchronized <T> T broadcastArray(T arrData, Operand operand, int from, int to, int rootRank) throws Mp4jException {
        CommUtils.isFromToLegal(from, to);
        CommUtils.isRangeInArray(arrData, to);
        checkRoot(rootRank);
        if (slaveNum == 1) return arrData;
        if (nativePath(operand, null) && isNativeArray(arrData)) {
            check(Native.hostBroadcast(handle, arrData, operand.dtype, from, to, rootRank));
            return arrData;
        }
        int[] froms = splitFroms(from, to), tos = splitTos(from, to);
        scatterArray(arrData, operand, froms, tos, rootRank);
        return allgatherArray(arrData, operand, froms, tos);
    }

    /** rank i receives root's [recvfroms[i], recvtos[i]) at the same offsets (PCS:900) */
    public synchronized <T> T scatterArray(T arrData, Operand operand, int[] recvfroms, int[] recvtos, int rootRank) throws Mp4jException {
        CommUtils.isfromsTosLegal(recvfroms, recvtos);
        CommUtils.isRangeInArray(arrData, CommUtils.maxEnd(recvtos));
        checkRoot(rootRank);
        if (slaveNum == 1) return arrData;
        if (nativePath(operand, null) && isNativeArray(arrData)) {
            check(Native.hostScatter(handle, arrData, operand.dtype, toLong(recvfroms), toLong(recvtos), rootRank));
            return arrData;
        }
        Object[] payloads = null;
        if (rank == rootRank) {
            payloads = new Object[slaveNum];
            for (int q = 0; q < slaveNum; q++) payloads[q] = ArrayOps.slice(arrData, recvfroms[q], recvtos[q]);
        }
        Object mine = treeScatter(payloads, rootRank, operand);
        if (rank != rootRank) ArrayOps.put(arrData, recvfroms[rank], mine);
        return arrData;
    }

    /** rank r ends with the fully reduced [from + sum(counts[&lt;r]), + counts[r]) (PCS:1174) */
    public synchronized <T> T reduceScatterArray(T arrData, Operand operand, IOperator operator, int from, int[] counts) throws Mp4jException {
        CommUtils.isFromCountsLegal(from, counts);
        CommUtils.isRangeInArray(arrData, CommUtils.endOfCounts(from, counts));
        if (slaveNum == 1) return arrData;
        if (nativePath(operand, operator) && isNativeArray(arrData)) {
            check(Native.hostReduceScatter(handle, arrData, operand.dtype, operator.nativeCode(), from, toLong(counts)));
            return arrData;
        }
        int[] froms = CommUtils.getFromsFromCount(from, counts, slaveNum), tos = CommUtils.getTosFromCount(from, counts, slaveNum);
        Object[] parts = new Object[slaveNum];
        for (int q = 0; q < slaveNum; q++) parts[q] = ArrayOps.slice(arrData, froms[q], tos[q]);
        ArrayOps.put(arrData, froms[rank], ringReduceScatter(parts, arrayMerger(operator), operand));
        return arrData;
    }

    /** = reduce-scatter + gather (PCS:1390-1416) */
    public synchronized <T> T reduceArray(T arrData, Operand operand, IOperator operator, int from, int to, int rootRank) throws Mp4jException {
        CommUtils.isFromToLegal(from, to);
        CommUtils.isRangeInArray(arrData, to);
        checkRoot(rootRank);
        if (slaveNum == 1) return arrData;
        if (nativePath(operand, operator) && isNativeArray(arrData)) {
            check(Native.hostReduce(handle, arrData, operand.dtype, operator.nativeCode(), from, to, rootRank));
            return arrData;
        }
        int[] froms = splitFroms(from, to), tos = splitTos(from, to);
        reduceScatterArray(arrData, operand, operator, from, counts(froms, tos));
        return gatherArray(arrData, operand, froms, tos, rootRank);
    }

    /** = reduce-scatter + allgather (PCS:1733-1763) */
    public synchronized <T> T allreduceArray(T arrData, Operand operand, IOperator operator, int from, int to) throws Mp4jException {
        CommUtils.isFromToLegal(from, to);
        CommUtils.isRangeInArray(arrData, to);
        if (slaveNum == 1) return arrData;
        if (nativePath(operand, operator) && isNativeArray(arrData)) {
            check(Native.hostAllreduce(handle, arrData, operand.dtype, operator.nativeCode(), from, to));
            return arrData;
        }
        int[] froms = splitFroms(from, to), tos = splitTos(from, to);
        reduceScatterArray(arrData, operand, operator, from, counts(froms, tos));
        return allgatherArray(arrData, operand, froms, tos);
    }

    /** small arrays: everybody's whole array meets, each rank folds them in rank order (PCS:1776-1799) */
    public synchronized <T> T allreduceArrayRpc(T arrData, Operand operand, IOperator operator) throws Mp4jException {
        if (slaveNum == 1) return arrData;
        int n = ArrayOps.length(arrData);
        if (nativePath(operand, operator) && isNativeArray(arrData)) {
            check(Native.hostAllreduceRpc(handle, arrData, operand.dtype, operator.nativeCode(), n));
            return arrData;
        }
        Object[] all = ringAllgather(ArrayOps.slice(arrData, 0, n), operand);
        Object acc = all[0];
        for (int q = 1; q < slaveNum; q++) ArrayOps.reduce(acc, all[q], 0, n, 0, operator);
        ArrayOps.put(arrData, 0, acc);
        return arrData;
    }

    // the balanced split of [from, to): first p-1 ranks take (to-from)/p, the last takes the rest (PCS:1741-1756)
    private int[] splitFroms(int from, int to) {
        int avg = (to - from) / slaveNum;
        int[] f = new int[slaveNum];
        for (int r = 0; r < slaveNum; r++) f[r] = from + r * avg;
        return f;
    }
    private int[] splitTos(int from, int to) {
        int avg = (to - from) / slaveNum;
        int[] t = new int[slaveNum];
        for (int r = 0; r < slaveNum; r++) t[r] = r == slaveNum - 1 ? to : from + (r + 1) * avg;
        return t;
    }
    private static int[] counts(int[] froms, int[] tos) {
        int[] c = new int[froms.length];
        for (int i = 0; i < c.length; i++) c[i] = tos[i] - froms[i];
        return c;
    }

    // ------------------------------------------------------------------ scalars: boxed into 1-element arrays (PCS:787-829)
    @SuppressWarnings("unchecked")
    public <T> T broadcast(T value, Operand operand, int rootRank) throws Mp4jException {
        if (slaveNum == 1) return value;
        Object box = ArrayOps.box(value);
        broadcastArray(box, operand, 0, 1, rootRank);
        return (T) ArrayOps.unbox(box);
    }
    @SuppressWarnings("unchecked")
    public <T> T reduce(T value, Operand operand, IOperator operator, int rootRank) throws Mp4jException {
        if (slaveNum == 1) return value;
        Object box = ArrayOps.box(value);
        reduceArray(box, operand, operator, 0, 1, rootRank);
        return (T) ArrayOps.unbox(box);
    }
    @SuppressWarnings("unchecked")
    public <T> T allreduce(T value, Operand operand, IOperator operator) throws Mp4jException {
        if (slaveNum == 1) return value;
        Object box = ArrayOps.box(value);
        allreduceArray(box, operand, operator, 0, 1);
        return (T) ArrayOps.unbox(box);
    }
    @SuppressWarnings("unchecked")
    public <T> T allreduceRpc(T value, Operand operand, IOperator operator) throws Mp4jException {
        if (slaveNum == 1) return value;
        Object box = ArrayOps.box(value);
        allreduceArrayRpc(box, operand, operator);
        return (T) ArrayOps.unbox(box);
    }

    // ------------------------------------------------------------------ maps: Map<String, V>
    /** key placement of allreduceMap / reduceMap / broadcastMap: String.hashCode() % slaveNum, negatives wrapped (PCS:2063-2071) */
    private int partition(String key) { int m = key.hashCode() % slaveNum; return m < 0 ? m + slaveNum : m; }

    private <V> List<Map<String, V>> partitionMap(Map<String, V> map) {
        List<Map<String, V>> parts = new ArrayList<Map<String, V>>();
        for (int i = 0; i < slaveNum; i++) parts.add(new HashMap<String, V>());
        for (Map.Entry<String, V> e : map.entrySet()) parts.get(partition(e.getKey())).put(e.getKey(), e.getValue());
        return parts;
    }

    /** root receives the union of all maps; duplicate keys: the higher rank's value wins (PCS:582, 477-481) */
    @SuppressWarnings("unchecked")
    public synchronized <V> Map<String, V> gatherMap(Map<String, V> mapData, Operand operand, int rootRank) throws Mp4jException {
        checkRoot(rootRank);
        if (slaveNum == 1) return mapData;
        Map<Integer, Object> acc = treeGather(mapData, rootRank, operand);
        Map<String, V> out = new HashMap<String, V>();
        if (rank != rootRank) return out;
        for (int q = 0; q < slaveNum; q++) out.putAll((Map<String, V>) acc.get(q));
        return out;
    }

    /** list of the p maps indexed by origin rank (PCS:666) */
    @SuppressWarnings("unchecked")
    public synchronized <V> List<Map<String, V>> allgatherMap(Map<String, V> mapData, Operand operand) throws Mp4jException {
        List<Map<String, V>> out = new ArrayList<Map<String, V>>();
        if (slaveNum == 1) {
            out.add(mapData);
            return out;
        }
        for (Object m : ringAllgather(mapData, operand)) out.add((Map<String, V>) m);
        return out;
    }

    /** root's map to everybody: hash-partition on the root, scatter, allgather, merge (PCS:842-875) */
    public synchronized <V> Map<String, V> broadcastMap(Map<String, V> mapData, Operand operand, int rootRank) throws Mp4jException {
        checkRoot(rootRank);
        if (slaveNum == 1) return mapData;
        List<Map<String, V>> parts = rank == rootRank ? partitionMap(mapData) : null;
        Map<String, V> mine = scatterMap(parts, operand, rootRank);
        Map<String, V> out = new HashMap<String, V>();
        for (Map<String, V> m : allgatherMap(mine, operand)) out.putAll(m);
        return out;
    }

    /** rank i receives mapDataList.get(i) of the root (PCS:949) */
    @SuppressWarnings("unchecked")
    public synchronized <V> Map<String, V> scatterMap(List<Map<String, V>> mapDataList, Operand operand, int rootRank) throws Mp4jException {
        checkRoot(rootRank);
        if (slaveNum == 1) return mapDataList.get(0);
        Object[] payloads = null;
        if (rank == rootRank) {
            if (mapDataList == null || mapDataList.size() != slaveNum) throw new Mp4jException("scatterMap needs one map per rank on the root");
            payloads = mapDataList.toArray();
        }
        return (Map<String, V>) treeScatter(payloads, rootRank, operand);
    }

    /** mapDataList.get(i) = this rank's contribution to rank i; returns the merged map this rank owns (PCS:1299) */
    @SuppressWarnings("unchecked")
    public synchronized <V> Map<String, V> reduceScatterMap(List<Map<String, V>> mapDataList, Operand operand, IOperator operator) throws Mp4jException {
        if (mapDataList == null || mapDataList.size() != slaveNum) throw new Mp4jException("reduceScatterMap needs one map per rank");
        if (slaveNum == 1) return mapDataList.get(0);
        return (Map<String, V>) ringReduceScatter(mapDataList.toArray(), mapMerger(operator), operand);
    }

    /** = hash-partition + reduce-scatter + gather; non-root ranks get an empty map (PCS:1490-1508) */
    public synchronized <V> Map<String, V> reduceMap(Map<String, V> mapData, Operand operand, IOperator operator, int rootRank) throws Mp4jException {
        checkRoot(rootRank);
        if (slaveNum == 1) return mapData;
        Map<String, V> mine = reduceScatterMap(partitionMap(mapData), operand, operator);
        return gatherMap(mine, operand, rootRank);
    }

    /** = hash-partition + reduce-scatter + allgather + union of the p disjoint maps (PCS:2053-2081) */
    public synchronized <V> Map<String, V> allreduceMap(Map<String, V> mapData, Operand operand, IOperator operator) throws Mp4jException {
        if (slaveNum == 1) return mapData;
        Map<String, V> mine = reduceScatterMap(partitionMap(mapData), operand, operator);
        Map<String, V> out = new HashMap<String, V>();
        for (Map<String, V> m : allgatherMap(mine, operand)) out.putAll(m);
        return out;
    }

    /** T maps per rank (used by ThreadCommSlave): rank i receives lists.get(i) of the root (PCS:999) */
    @SuppressWarnings("unchecked")
    protected synchronized <V> List<Map<String, V>> scatterMapSpecial(List<List<Map<String, V>>> lists, Operand operand, int rootRank) throws Mp4jException {
        checkRoot(rootRank);
        if (slaveNum == 1) return lists.get(0);
        Object[] payloads = null;
        if (rank == rootRank) {
            if (lists == null || lists.size() != slaveNum) throw new Mp4jException("scatterMap needs [slaveNum][threadNum] maps on the root");
            payloads = lists.toArray();
        }
        return (List<Map<String, V>>) treeScatter(payloads, rootRank, operand);
    }

    /** T maps per rank: lists.get(i).get(t) = contribution to thread t of rank i; returns this rank's T merged maps (PCS:1212) */
    @SuppressWarnings("unchecked")
    protected synchronized <V> List<Map<String, V>> reduceScatterMapSpecial(List<List<Map<String, V>>> lists, Operand operand, final IOperator operator) throws Mp4jException {
        if (lists == null || lists.size() != slaveNum) throw new Mp4jException("reduceScatterMap needs [slaveNum][threadNum] maps");
        if (slaveNum == 1) return lists.get(0);
        Merger merge = new Merger() {
            @Override public Object merge(Object local, Object incoming) throws Mp4jException {
                List<Map<String, V>> l = (List<Map<String, V>>) local, in = (List<Map<String, V>>) incoming;
                for (int t = 0; t < l.size(); t++) ArrayOps.mergeMap(l.get(t), in.get(t), operator);
                return l;
            }
        };
        return (List<Map<String, V>>) ringReduceScatter(lists.toArray(), merge, operand);
    }

    // ------------------------------------------------------------------ set / list helpers (PCS:1583-1718, 2099-2211)
    // reduceMap / allreduceMap over an Object operand with addAll / retainAll / append operators; the
    // scalar forms wrap the value in a one-entry map under the key "key", as the reference does.
    static <E> IObjectOperator<Set<E>> setUnion() {
        return new IObjectOperator<Set<E>>() {
            @Override public Set<E> apply(Set<E> a, Set<E> b) { a.addAll(b); return a; }
        };
    }
    static <E> IObjectOperator<Set<E>> setIntersection() {
        return new IObjectOperator<Set<E>>() {
            @Override public Set<E> apply(Set<E> a, Set<E> b) { a.retainAll(b); return a; }
        };
    }
    static <E> IObjectOperator<List<E>> listConcat() {
        return new IObjectOperator<List<E>>() {
            @Override public List<E> apply(List<E> a, List<E> b) { a.addAll(b); return a; }
        };
    }
    /** containers of Serializable elements travel with java.io serialization */
    static Operand containerOperand() { return new ObjectOperand<Object>(null, Object.class, false); }
    static <X> Map<String, X> single(X value) {
        Map<String, X> m = new HashMap<String, X>();
        m.put("key", value);
        return m;
    }

    public <E> Map<String, Set<E>> reduceMapSetUnion(Map<String, Set<E>> mapData, int rootRank) throws Mp4jException {
        return reduceMap(mapData, containerOperand(), ProcessCommSlave.<E>setUnion(), rootRank);
    }
    public <E> Set<E> reduceSetUnion(Set<E> setData, int rootRank) throws Mp4jException {
        return reduceMapSetUnion(single(setData), rootRank).get("key");
    }
    public <E> Map<String, Set<E>> reduceMapSetIntersection(Map<String, Set<E>> mapData, int rootRank) throws Mp4jException {
        return reduceMap(mapData, containerOperand(), ProcessCommSlave.<E>setIntersection(), rootRank);
    }
    public <E> Set<E> reduceSetIntersection(Set<E> setData, int rootRank) throws Mp4jException {
        return reduceMapSetIntersection(single(setData), rootRank).get("key");
    }
    public <E> Map<String, List<E>> reduceMapListConcat(Map<String, List<E>> mapData, int rootRank) throws Mp4jException {
        return reduceMap(mapData, containerOperand(), ProcessCommSlave.<E>listConcat(), rootRank);
    }
    public <E> List<E> reduceListConcat(List<E> listData, int rootRank) throws Mp4jException {
        return reduceMapListConcat(single(listData), rootRank).get("key");
    }
    public <E> Map<String, Set<E>> allreduceMapSetUnion(Map<String, Set<E>> mapData) throws Mp4jException {
        return allreduceMap(mapData, containerOperand(), ProcessCommSlave.<E>setUnion());
    }
    public <E> Set<E> allreduceSetUnion(Set<E> setData) throws Mp4jException {
        return allreduceMapSetUnion(single(setData)).get("key");
    }
    public <E> Map<String, Set<E>> allreduceMapSetIntersection(Map<String, Set<E>> mapData) throws Mp4jException {
        return allreduceMap(mapData, containerOperand(), ProcessCommSlave.<E>setIntersection());
    }
    public <E> Set<E> allreduceSetIntersection(Set<E> setData) throws Mp4jException {
        return allreduceMapSetIntersection(single(setData)).get("key");
    }
    public <E> Map<String, List<E>> allreduceMapListConcat(Map<String, List<E>> mapData) throws Mp4jException {
        return allreduceMap(mapData, containerOperand(), ProcessCommSlave.<E>listConcat());
    }
    public <E> List<E> allreduceListConcat(List<E> listData) throws Mp4jException {
        return allreduceMapListConcat(single(listData)).get("key");
    }
    /** convenience for callers that start from arrays */
    public static <E> Set<E> newSet(java.util.Collection<E> c) { return new HashSet<E>(c); }

    // ------------------------------------------------------------------ device arrays (raw CUDA addresses): the sm_100a kernels
    public synchronized void initGpu(int device, long heapBytes, long stagingBytes) throws Mp4jException {
        gpu = Native.gpuInit(handle, device, heapBytes, stagingBytes);
        if (gpu == 0) throw new Mp4jException(Native.lastError());
    }
    /** handle of the GPU plane for the thread-level device methods of ThreadCommSlave (0 before initGpu) */
    long gpuHandle() { return gpu; }
    static int deviceOp(IOperator op) throws Mp4jException {
        if (op.nativeCode() < 0) throw new Mp4jException("device collectives take the predefined operators of Operators.*");
        return op.nativeCode();
    }
    public void allreduceDevice(long devPtr, Operand operand, IOperator operator, long from, long to, long stream) throws Mp4jException {
        check(Native.gpuAllreduce(gpu, devPtr, operand.dtype, deviceOp(operator), from, to, 0, stream));
    }
    public void reduceScatterDevice(long devPtr, Operand operand, IOperator operator, long from, long[] counts, long stream) throws Mp4jException {
        check(Native.gpuReduceScatter(gpu, devPtr, operand.dtype, deviceOp(operator), from, counts, stream));
    }
    public void allgatherDevice(long devPtr, Operand operand, long[] froms, long[] tos, long stream) throws Mp4jException {
        check(Native.gpuAllgather(gpu, devPtr, operand.dtype, froms, tos, stream));
    }
    public void broadcastDevice(long devPtr, Operand operand, long from, long to, int root, long stream) throws Mp4jException {
        check(Native.gpuBroadcast(gpu, devPtr, operand.dtype, from, to, root, stream));
    }
}
