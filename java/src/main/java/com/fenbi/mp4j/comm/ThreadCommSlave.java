package com.fenbi.mp4j.comm;

import com.fenbi.mp4j.exception.Mp4jException;
import com.fenbi.mp4j.operand.Operand;
import com.fenbi.mp4j.operator.IOperator;
import com.fenbi.mp4j.utils.ArrayOps;
import com.fenbi.mp4j.utils.CommUtils;

import java.util.ArrayList;
import java.util.HashMap;
import java.util.List;
import java.util.Map;
import java.util.Set;
import java.util.concurrent.CyclicBarrier;

/** Process x thread communicator with the reference's method names (J/comm/ThreadCommSlave.java:75-2455):
 *  one instance per process shared by threadNum threads, each of which calls setThreadId(t) once.
 *
 *  Every collective exists twice: xxx(...) - all slaveNum x threadNum threads take part, ranges are
 *  int[slaveNum][threadNum], roots are (rootRank, rootThreadId) - and xxxProcess(...) - a pass-through
 *  to the process communicator for "one thread per process calls this".
 *
 *  Thread level = delegate pattern: every thread deposits its argument, the delegate thread (the root
 *  thread on the root process, thread 0 elsewhere: TCS:486) combines the T deposits (thread-id order),
 *  runs the process collective and fans the result out; two thread barriers per call plus one that
 *  protects the deposit slots. On CUDA tensors the same contract is ONE fused kernel
 *  (csrc/gpu/thread_fused_ws.cu); this class is the host-array form. Map results are shared by the
 *  threads of a process, as in the reference (README warning, TCS:2214-2217): clone before mutating. */
public class ThreadCommSlave {
    private final ProcessCommSlave process;
    private final int threadNum;
    private final CyclicBarrier barrier;
    private final ThreadLocal<Integer> threadId = new ThreadLocal<Integer>() {
        @Override protected Integer initialValue() { return 0; }
    };
    private final Object[] slots;
    private volatile Object shared;
    private volatile Throwable failure;

    public ThreadCommSlave(String loginName, int threadNum, String masterHost, int masterPort) throws Mp4jException {
        if (threadNum < 1) throw new Mp4jException("threadNum must be >= 1");
        this.process = new ProcessCommSlave(loginName, masterHost, masterPort);
        this.threadNum = threadNum;
        this.barrier = new CyclicBarrier(threadNum);
        this.slots = new Object[threadNum];
    }

    public int getThreadNum() { return threadNum; }
    public boolean isClosed() { return process.isClosed(); }
    public void setThreadId(int id) throws Mp4jException {
        if (id < 0 || id >= threadNum) throw new Mp4jException("threadId must be in [0, threadNum)");
        threadId.set(id);
    }
    public int getThreadId() { return threadId.get(); }
    public int getRank() { return process.getRank(); }
    public int getSlaveNum() { return process.getSlaveNum(); }

    public void threadBarrier() throws Mp4jException {
        if (threadNum == 1) return;
        try {
            barrier.await();
        } catch (Exception e) {
            throw new Mp4jException(e);
        }
    }
    /** thread barrier, thread 0 takes the master barrier, thread barrier (TCS:151) */
    public void barrier() throws Mp4jException {
        threadBarrier();
        if (getThreadId() == 0) process.barrier();
        threadBarrier();
    }

    // default: only thread 0 of rank 0 speaks; otherwise the message is prefixed with the thread id (TCS:180-257)
    public void info(String info) throws Mp4jException { info(info, true); }
    public void info(String info, boolean onlyRank0Thread0) throws Mp4jException {
        if (onlyRank0Thread0) {
            if (getRank() == 0 && getThreadId() == 0) process.info(info, true);
        } else {
            process.info("[threadId=" + getThreadId() + "] " + info, false);
        }
    }
    public void debug(String debug) throws Mp4jException { debug(debug, true); }
    public void debug(String debug, boolean onlyRank0Thread0) throws Mp4jException {
        if (onlyRank0Thread0) {
            if (getRank() == 0 && getThreadId() == 0) process.debug(debug, true);
        } else {
            process.debug("[threadId=" + getThreadId() + "] " + debug, false);
        }
    }
    public void error(String error) throws Mp4jException { process.error("[threadId=" + getThreadId() + "] " + error); }
    public void exception(Exception e) throws Mp4jException { process.exception(e); }
    public void close(int code) throws Mp4jException { process.close(code); }

    // ------------------------------------------------------------------ delegate machinery
    private interface Body { Object run(Object[] deposits) throws Mp4jException; }

    /** All T threads call with their own deposit; thread `delegateTid` runs body over the deposits (indexed
     *  by thread id) and its result reaches every thread. A failure on the delegate surfaces on every thread. */
    private Object delegate(Object mine, int delegateTid, Body body) throws Mp4jException {
        int tid = getThreadId();
        slots[tid] = mine;
        threadBarrier();
        if (tid == delegateTid) {
            try {
                shared = body.run(slots.clone());
                failure = null;
            } catch (Mp4jException e) {
                shared = null;
                failure = e;
            } catch (RuntimeException e) {
                shared = null;
                failure = e;
            }
        }
        threadBarrier();
        Object result = shared;
        Throwable f = failure;
        threadBarrier();  // everybody has read the result before the next call may overwrite it
        if (f != null) throw (f instanceof Mp4jException) ? (Mp4jException) f : new Mp4jException(f);
        return result;
    }

    /** only the root process honours rootThreadId; every other process delegates through thread 0 (TCS:486) */
    private int realRoot(int rootRank, int rootThreadId) throws Mp4jException {
        if (rootThreadId < 0 || rootThreadId >= threadNum) throw new Mp4jException("rootThreadId out of range");
        return getRank() == rootRank ? rootThreadId : 0;
    }

    /** arrs[dst][from, to) = op over the threads, dst first, then ascending thread id (DOP:441-457) */
    private static void localReduce(Object[] arrs, IOperator op, int from, int to, int dst) throws Mp4jException {
        for (int t = 0; t < arrs.length; t++)
            if (t != dst) ArrayOps.reduce(arrs[dst], arrs[t], from, to, 0, op);
    }

    private static int[] rowSums(int[][] counts) {
        int[] out = new int[counts.length];
        for (int r = 0; r < counts.length; r++) for (int c : counts[r]) out[r] += c;
        return out;
    }
    private static int[] firstColumn(int[][] counts) {
        int[] out = new int[counts.length];
        for (int r = 0; r < counts.length; r++) out[r] = counts[r][0];
        return out;
    }

    // ------------------------------------------------------------------ arrays, all threads (2-D ranges)
    /** the result lands in the root thread's array (TCS:448) */
    public <T> T gatherArray(final T arrData, final Operand operand, final int[][] sendfroms, final int[][] sendtos,
                             final int rootRank, int rootThreadId) throws Mp4jException {
        final int[] pf = CommUtils.getProcessFroms(sendfroms), pt = CommUtils.getProcessTos(sendtos);
        if (threadNum == 1) return process.gatherArray(arrData, operand, pf, pt, rootRank);
        CommUtils.isfromsTosLegal(sendfroms, sendtos, threadNum);
        CommUtils.isRangeInArray(arrData, CommUtils.maxEnd(sendtos));
        final int rr = realRoot(rootRank, rootThreadId), r = getRank();
        delegate(arrData, rr, new Body() {
            @Override public Object run(Object[] arrs) throws Mp4jException {
                for (int t = 0; t < threadNum; t++) ArrayOps.copy(arrs[t], arrs[rr], sendfroms[r][t], sendtos[r][t]);
                process.gatherArray(arrs[rr], operand, pf, pt, rootRank);
                return null;
            }
        });
        return arrData;
    }

    /** every thread's array ends with all slaveNum x threadNum segments (TCS:621, full copy-out TCS:676) */
    public <T> T allgatherArray(final T arrData, final Operand operand, final int[][] froms, final int[][] tos) throws Mp4jException {
        final int[] pf = CommUtils.getProcessFroms(froms), pt = CommUtils.getProcessTos(tos);
        if (threadNum == 1) return process.allgatherArray(arrData, operand, pf, pt);
        CommUtils.isfromsTosLegal(froms, tos, threadNum);
        CommUtils.isRangeInArray(arrData, CommUtils.maxEnd(tos));
        final int r = getRank();
        delegate(arrData, 0, new Body() {
            @Override public Object run(Object[] arrs) throws Mp4jException {
                for (int t = 1; t < threadNum; t++) ArrayOps.copy(arrs[t], arrs[0], froms[r][t], tos[r][t]);
                process.allgatherArray(arrs[0], operand, pf, pt);
                for (int t = 1; t < threadNum; t++)
                    for (int q = 0; q < pf.length; q++) ArrayOps.copy(arrs[0], arrs[t], pf[q], pt[q]);
                return null;
            }
        });
        return arrData;
    }

    public <T> T broadcastArray(final T arrData, final Operand operand, final int from, final int to, final int rootRank,
                                int rootThreadId) throws Mp4jException {
        if (threadNum == 1) return process.broadcastArray(arrData, operand, from, to, rootRank);
        CommUtils.isFromToLegal(from, to);
        CommUtils.isRangeInArray(arrData, to);
        final int rr = realRoot(rootRank, rootThreadId);
        delegate(arrData, rr, new Body() {
            @Override public Object run(Object[] arrs) throws Mp4jException {
                process.broadcastArray(arrs[rr], operand, from, to, rootRank);
                for (int t = 0; t < threadNum; t++) ArrayOps.copy(arrs[rr], arrs[t], from, to);
                return null;
            }
        });
        return arrData;
    }

    /** thread (r, t) receives the root thread's [recvfroms[r][t], recvtos[r][t]) (TCS:1010) */
    public <T> T scatterArray(final T arrData, final Operand operand, final int[][] recvfroms, final int[][] recvtos,
                              final int rootRank, int rootThreadId) throws Mp4jException {
        final int[] pf = CommUtils.getProcessFroms(recvfroms), pt = CommUtils.getProcessTos(recvtos);
        if (threadNum == 1) return process.scatterArray(arrData, operand, pf, pt, rootRank);
        CommUtils.isfromsTosLegal(recvfroms, recvtos, threadNum);
        CommUtils.isRangeInArray(arrData, CommUtils.maxEnd(recvtos));
        final int rr = realRoot(rootRank, rootThreadId), r = getRank();
        delegate(arrData, rr, new Body() {
            @Override public Object run(Object[] arrs) throws Mp4jException {
                process.scatterArray(arrs[rr], operand, pf, pt, rootRank);
                for (int t = 0; t < threadNum; t++) ArrayOps.copy(arrs[rr], arrs[t], recvfroms[r][t], recvtos[r][t]);
                return null;
            }
        });
        return arrData;
    }

    /** counts is [slaveNum][threadNum]; thread (r, t) ends with its fully reduced slice (TCS:1174) */
    public <T> T reduceScatterArray(final T arrData, final Operand operand, final IOperator operator, final int from,
                                    final int[][] counts) throws Mp4jException {
        if (threadNum == 1) return process.reduceScatterArray(arrData, operand, operator, from, firstColumn(counts));
        CommUtils.isFromCountsLegal(from, counts);
        CommUtils.isRangeInArray(arrData, CommUtils.endOfCounts(from, counts));
        final int r = getRank();
        final int[] pcounts = rowSums(counts);
        int total = 0;
        for (int c : pcounts) total += c;
        final int totalTo = from + total;
        delegate(arrData, 0, new Body() {
            @Override public Object run(Object[] arrs) throws Mp4jException {
                localReduce(arrs, operator, from, totalTo, 0);
                process.reduceScatterArray(arrs[0], operand, operator, from, pcounts);
                int off = from;
                for (int q = 0; q < r; q++) off += pcounts[q];
                for (int t = 0; t < threadNum; t++) {
                    ArrayOps.copy(arrs[0], arrs[t], off, off + counts[r][t]);
                    off += counts[r][t];
                }
                return null;
            }
        });
        return arrData;
    }

    /** the root thread's array holds the reduction; other threads keep their input (TCS:1396, 1433) */
    public <T> T reduceArray(final T arrData, final Operand operand, final IOperator operator, final int from, final int to,
                             final int rootRank, int rootThreadId) throws Mp4jException {
        if (threadNum == 1) return process.reduceArray(arrData, operand, operator, from, to, rootRank);
        CommUtils.isFromToLegal(from, to);
        CommUtils.isRangeInArray(arrData, to);
        final int rr = realRoot(rootRank, rootThreadId);
        delegate(arrData, rr, new Body() {
            @Override public Object run(Object[] arrs) throws Mp4jException {
                localReduce(arrs, operator, from, to, rr);
                process.reduceArray(arrs[rr], operand, operator, from, to, rootRank);
                return null;
            }
        });
        return arrData;
    }

    /** all slaveNum x threadNum arrays are reduced; every thread's array holds the result (TCS:1915-1963) */
    public <T> T allreduceArray(final T arrData, final Operand operand, final IOperator operator, final int from, final int to) throws Mp4jException {
        if (threadNum == 1) return process.allreduceArray(arrData, operand, operator, from, to);
        CommUtils.isFromToLegal(from, to);
        CommUtils.isRangeInArray(arrData, to);
        delegate(arrData, 0, new Body() {
            @Override public Object run(Object[] arrs) throws Mp4jException {
                localReduce(arrs, operator, from, to, 0);
                process.allreduceArray(arrs[0], operand, operator, from, to);
                for (int t = 1; t < threadNum; t++) ArrayOps.copy(arrs[0], arrs[t], from, to);
                return null;
            }
        });
        return arrData;
    }

    public <T> T allreduceArrayRpc(final T arrData, final Operand operand, final IOperator operator) throws Mp4jException {
        if (threadNum == 1) return process.allreduceArrayRpc(arrData, operand, operator);
        final int n = ArrayOps.length(arrData);
        delegate(arrData, 0, new Body() {
            @Override public Object run(Object[] arrs) throws Mp4jException {
                localReduce(arrs, operator, 0, n, 0);
                process.allreduceArrayRpc(arrs[0], operand, operator);
                for (int t = 1; t < threadNum; t++) ArrayOps.copy(arrs[0], arrs[t], 0, n);
                return null;
            }
        });
        return arrData;
    }

    // ------------------------------------------------------------------ scalars, all threads
    private static Object threadFold(Object[] vals, IOperator op, int first) throws Mp4jException {
        Object acc = vals[first];
        for (int t = 0; t < vals.length; t++) if (t != first) acc = ArrayOps.apply(op, acc, vals[t]);
        return acc;
    }

    @SuppressWarnings("unchecked")
    public <T> T broadcast(T value, final Operand operand, final int rootRank, int rootThreadId) throws Mp4jException {
        if (threadNum == 1) return process.broadcast(value, operand, rootRank);
        final int rr = realRoot(rootRank, rootThreadId);
        return (T) delegate(value, rr, new Body() {
            @Override public Object run(Object[] vals) throws Mp4jException { return process.broadcast(vals[rr], operand, rootRank); }
        });
    }

    /** the root thread gets the reduction; every other thread gets its own input back (TCS:1433) */
    @SuppressWarnings("unchecked")
    public <T> T reduce(T value, final Operand operand, final IOperator operator, final int rootRank, int rootThreadId) throws Mp4jException {
        if (threadNum == 1) return process.reduce(value, operand, operator, rootRank);
        final int rr = realRoot(rootRank, rootThreadId);
        Object res = delegate(value, rr, new Body() {
            @Override public Object run(Object[] vals) throws Mp4jException {
                return process.reduce(threadFold(vals, operator, rr), operand, operator, rootRank);
            }
        });
        return (getRank() == rootRank && getThreadId() == rr) ? (T) res : value;
    }

    @SuppressWarnings("unchecked")
    public <T> T allreduce(T value, final Operand operand, final IOperator operator) throws Mp4jException {
        if (threadNum == 1) return process.allreduce(value, operand, operator);
        return (T) delegate(value, 0, new Body() {
            @Override public Object run(Object[] vals) throws Mp4jException {
                return process.allreduce(threadFold(vals, operator, 0), operand, operator);
            }
        });
    }

    @SuppressWarnings("unchecked")
    public <T> T allreduceRpc(T value, final Operand operand, final IOperator operator) throws Mp4jException {
        if (threadNum == 1) return process.allreduceRpc(value, operand, operator);
        return (T) delegate(value, 0, new Body() {
            @Override public Object run(Object[] vals) throws Mp4jException {
                return process.allreduceRpc(threadFold(vals, operator, 0), operand, operator);
            }
        });
    }

    // ------------------------------------------------------------------ maps, all threads
    /** union (operator == null: later thread wins) or operator-merge of the T thread maps, thread-id order */
    @SuppressWarnings("unchecked")
    private static <V> Map<String, V> mergeThreadMaps(Object[] maps, IOperator op) throws Mp4jException {
        Map<String, V> out = new HashMap<String, V>();
        for (Object m : maps) {
            if (m == null) continue;
            if (op == null) out.putAll((Map<String, V>) m);
            else ArrayOps.mergeMap(out, (Map<String, V>) m, op);
        }
        return out;
    }

    /** root thread gets the union of all maps, everybody else null (TCS:544, 1583) */
    @SuppressWarnings("unchecked")
    public <V> Map<String, V> gatherMap(Map<String, V> mapData, final Operand operand, final int rootRank, int rootThreadId) throws Mp4jException {
        if (threadNum == 1) return process.gatherMap(mapData, operand, rootRank);
        final int rr = realRoot(rootRank, rootThreadId);
        Object res = delegate(mapData, rr, new Body() {
            @Override public Object run(Object[] maps) throws Mp4jException {
                return process.gatherMap(ThreadCommSlave.<V>mergeThreadMaps(maps, null), operand, rootRank);
            }
        });
        return (getRank() == rootRank && getThreadId() == rr) ? (Map<String, V>) res : null;
    }

    /** list indexed by origin rank of the (thread-merged) maps; shared by the threads of a process (TCS:717) */
    @SuppressWarnings("unchecked")
    public <V> List<Map<String, V>> allgatherMap(Map<String, V> mapData, final Operand operand) throws Mp4jException {
        if (threadNum == 1) return process.allgatherMap(mapData, operand);
        return (List<Map<String, V>>) delegate(mapData, 0, new Body() {
            @Override public Object run(Object[] maps) throws Mp4jException {
                return process.allgatherMap(ThreadCommSlave.<V>mergeThreadMaps(maps, null), operand);
            }
        });
    }

    @SuppressWarnings("unchecked")
    public <V> Map<String, V> broadcastMap(Map<String, V> mapData, final Operand operand, final int rootRank, int rootThreadId) throws Mp4jException {
        if (threadNum == 1) return process.broadcastMap(mapData, operand, rootRank);
        final int rr = realRoot(rootRank, rootThreadId);
        return (Map<String, V>) delegate(mapData, rr, new Body() {
            @Override public Object run(Object[] maps) throws Mp4jException {
                return process.broadcastMap((Map<String, V>) maps[rr], operand, rootRank);
            }
        });
    }

    /** the root thread holds [slaveNum][threadNum] maps; thread (r, t) receives map [r][t] (TCS:1097) */
    @SuppressWarnings("unchecked")
    public <V> Map<String, V> scatterMap(List<List<Map<String, V>>> mapDataListList, final Operand operand, final int rootRank,
                                         int rootThreadId) throws Mp4jException {
        if (threadNum == 1) {
            List<Map<String, V>> firsts = null;
            if (mapDataListList != null) {
                firsts = new ArrayList<Map<String, V>>();
                for (List<Map<String, V>> l : mapDataListList) firsts.add(l.get(0));
            }
            return process.scatterMap(firsts, operand, rootRank);
        }
        final int rr = realRoot(rootRank, rootThreadId);
        List<Map<String, V>> mine = (List<Map<String, V>>) delegate(mapDataListList, rr, new Body() {
            @Override public Object run(Object[] lists) throws Mp4jException {
                return process.scatterMapSpecial((List<List<Map<String, V>>>) lists[rr], operand, rootRank);
            }
        });
        return mine.get(getThreadId());
    }

    /** every thread contributes [slaveNum][threadNum] maps; thread (r, t) gets the reduced map [r][t] (TCS:1286) */
    @SuppressWarnings("unchecked")
    public <V> Map<String, V> reduceScatterMap(List<List<Map<String, V>>> mapDataListList, final Operand operand,
                                               final IOperator operator) throws Mp4jException {
        if (threadNum == 1) {
            List<Map<String, V>> firsts = new ArrayList<Map<String, V>>();
            for (List<Map<String, V>> l : mapDataListList) firsts.add(l.get(0));
            return process.reduceScatterMap(firsts, operand, operator);
        }
        final int p = getSlaveNum();
        List<Map<String, V>> mine = (List<Map<String, V>>) delegate(mapDataListList, 0, new Body() {
            @Override public Object run(Object[] deposits) throws Mp4jException {
                List<List<Map<String, V>>> merged = new ArrayList<List<Map<String, V>>>();
                for (int q = 0; q < p; q++) {
                    List<Map<String, V>> row = new ArrayList<Map<String, V>>();
                    for (int t = 0; t < threadNum; t++) {
                        Object[] contributions = new Object[threadNum];
                        for (int s = 0; s < threadNum; s++) contributions[s] = ((List<List<Map<String, V>>>) deposits[s]).get(q).get(t);
                        row.add(ThreadCommSlave.<V>mergeThreadMaps(contributions, operator));
                    }
                    merged.add(row);
                }
                return process.reduceScatterMapSpecial(merged, operand, operator);
            }
        });
        return mine.get(getThreadId());
    }

    /** root thread gets the reduced map, everybody else null (TCS:1549, 1583) */
    @SuppressWarnings("unchecked")
    public <V> Map<String, V> reduceMap(Map<String, V> mapData, final Operand operand, final IOperator operator, final int rootRank,
                                        int rootThreadId) throws Mp4jException {
        if (threadNum == 1) return process.reduceMap(mapData, operand, operator, rootRank);
        final int rr = realRoot(rootRank, rootThreadId);
        Object res = delegate(mapData, rr, new Body() {
            @Override public Object run(Object[] maps) throws Mp4jException {
                return process.reduceMap(ThreadCommSlave.<V>mergeThreadMaps(maps, operator), operand, operator, rootRank);
            }
        });
        return (getRank() == rootRank && getThreadId() == rr) ? (Map<String, V>) res : null;
    }

    /** the result map is SHARED by the threads of a process (TCS:2179, 2214-2217) */
    @SuppressWarnings("unchecked")
    public <V> Map<String, V> allreduceMap(Map<String, V> mapData, final Operand operand, final IOperator operator) throws Mp4jException {
        if (threadNum == 1) return process.allreduceMap(mapData, operand, operator);
        return (Map<String, V>) delegate(mapData, 0, new Body() {
            @Override public Object run(Object[] maps) throws Mp4jException {
                return process.allreduceMap(ThreadCommSlave.<V>mergeThreadMaps(maps, operator), operand, operator);
            }
        });
    }

    // ------------------------------------------------------------------ set / list helpers, all threads (TCS:1607-1863, 2241-2455)
    public <E> Map<String, Set<E>> reduceMapSetUnion(Map<String, Set<E>> mapData, int rootRank, int rootThreadId) throws Mp4jException {
        return reduceMap(mapData, ProcessCommSlave.containerOperand(), ProcessCommSlave.<E>setUnion(), rootRank, rootThreadId);
    }
    public <E> Set<E> reduceSetUnion(Set<E> setData, int rootRank, int rootThreadId) throws Mp4jException {
        Map<String, Set<E>> m = reduceMapSetUnion(ProcessCommSlave.single(setData), rootRank, rootThreadId);
        return m == null ? null : m.get("key");
    }
    public <E> Map<String, Set<E>> reduceMapSetIntersection(Map<String, Set<E>> mapData, int rootRank, int rootThreadId) throws Mp4jException {
        return reduceMap(mapData, ProcessCommSlave.containerOperand(), ProcessCommSlave.<E>setIntersection(), rootRank, rootThreadId);
    }
    public <E> Set<E> reduceSetIntersection(Set<E> setData, int rootRank, int rootThreadId) throws Mp4jException {
        Map<String, Set<E>> m = reduceMapSetIntersection(ProcessCommSlave.single(setData), rootRank, rootThreadId);
        return m == null ? null : m.get("key");
    }
    public <E> Map<String, List<E>> reduceMapListConcat(Map<String, List<E>> mapData, int rootRank, int rootThreadId) throws Mp4jException {
        return reduceMap(mapData, ProcessCommSlave.containerOperand(), ProcessCommSlave.<E>listConcat(), rootRank, rootThreadId);
    }
    public <E> List<E> reduceListConcat(List<E> listData, int rootRank, int rootThreadId) throws Mp4jException {
        Map<String, List<E>> m = reduceMapListConcat(ProcessCommSlave.single(listData), rootRank, rootThreadId);
        return m == null ? null : m.get("key");
    }
    public <E> Map<String, Set<E>> allreduceMapSetUnion(Map<String, Set<E>> mapData) throws Mp4jException {
        return allreduceMap(mapData, ProcessCommSlave.containerOperand(), ProcessCommSlave.<E>setUnion());
    }
    public <E> Set<E> allreduceSetUnion(Set<E> setData) throws Mp4jException {
        return allreduceMapSetUnion(ProcessCommSlave.single(setData)).get("key");
    }
    public <E> Map<String, Set<E>> allreduceMapSetIntersection(Map<String, Set<E>> mapData) throws Mp4jException {
        return allreduceMap(mapData, ProcessCommSlave.containerOperand(), ProcessCommSlave.<E>setIntersection());
    }
    public <E> Set<E> allreduceSetIntersection(Set<E> setData) throws Mp4jException {
        return allreduceMapSetIntersection(ProcessCommSlave.single(setData)).get("key");
    }
    public <E> Map<String, List<E>> allreduceMapListConcat(Map<String, List<E>> mapData) throws Mp4jException {
        return allreduceMap(mapData, ProcessCommSlave.containerOperand(), ProcessCommSlave.<E>listConcat());
    }
    public <E> List<E> allreduceListConcat(List<E> listData) throws Mp4jException {
        return allreduceMapListConcat(ProcessCommSlave.single(listData)).get("key");
    }

    // ------------------------------------------------------------------ xxxProcess: one thread per process calls these
    public <T> T gatherArrayProcess(T arrData, Operand operand, int[] sendfroms, int[] sendtos, int rootRank) throws Mp4jException {
        return process.gatherArray(arrData, operand, sendfroms, sendtos, rootRank);
    }
    public <V> Map<String, V> gatherMapProcess(Map<String, V> mapData, Operand operand, int rootRank) throws Mp4jException {
        return process.gatherMap(mapData, operand, rootRank);
    }
    public <T> T allgatherArrayProcess(T arrData, Operand operand, int[] froms, int[] tos) throws Mp4jException {
        return process.allgatherArray(arrData, operand, froms, tos);
    }
    public <V> List<Map<String, V>> allgatherMapProcess(Map<String, V> mapData, Operand operand) throws Mp4jException {
        return process.allgatherMap(mapData, operand);
    }
    public <T> T broadcastArrayProcess(T arrData, Operand operand, int from, int to, int rootRank) throws Mp4jException {
        return process.broadcastArray(arrData, operand, from, to, rootRank);
    }
    public <T> T broadcastProcess(T value, Operand operand, int rootRank) throws Mp4jException {
        return process.broadcast(value, operand, rootRank);
    }
    public <V> Map<String, V> broadcastMapProcess(Map<String, V> mapData, Operand operand, int rootRank) throws Mp4jException {
        return process.broadcastMap(mapData, operand, rootRank);
    }
    public <T> T scatterArrayProcess(T arrData, Operand operand, int[] recvfroms, int[] recvtos, int rootRank) throws Mp4jException {
        return process.scatterArray(arrData, operand, recvfroms, recvtos, rootRank);
    }
    public <V> Map<String, V> scatterMapProcess(List<Map<String, V>> mapDataList, Operand operand, int rootRank) throws Mp4jException {
        return process.scatterMap(mapDataList, operand, rootRank);
    }
    public <T> T reduceScatterArrayProcess(T arrData, Operand operand, IOperator operator, int from, int[] counts) throws Mp4jException {
        return process.reduceScatterArray(arrData, operand, operator, from, counts);
    }
    public <V> Map<String, V> reduceScatterMapProcess(List<Map<String, V>> mapDataList, Operand operand, IOperator operator) throws Mp4jException {
        return process.reduceScatterMap(mapDataList, operand, operator);
    }
    public <T> T reduceArrayProcess(T arrData, Operand operand, IOperator operator, int from, int to, int rootRank) throws Mp4jException {
        return process.reduceArray(arrData, operand, operator, from, to, rootRank);
    }
    public <T> T reduceProcess(T value, Operand operand, IOperator operator, int rootRank) throws Mp4jException {
        return process.reduce(value, operand, operator, rootRank);
    }
    public <V> Map<String, V> reduceMapProcess(Map<String, V> mapData, Operand operand, IOperator operator, int rootRank) throws Mp4jException {
        return process.reduceMap(mapData, operand, operator, rootRank);
    }
    public <E> Map<String, Set<E>> reduceMapSetUnionProcess(Map<String, Set<E>> mapData, int rootRank) throws Mp4jException {
        return process.reduceMapSetUnion(mapData, rootRank);
    }
    public <E> Set<E> reduceSetUnionProcess(Set<E> setData, int rootRank) throws Mp4jException {
        return process.reduceSetUnion(setData, rootRank);
    }
    public <E> Map<String, Set<E>> reduceMapSetIntersectionProcess(Map<String, Set<E>> mapData, int rootRank) throws Mp4jException {
        return process.reduceMapSetIntersection(mapData, rootRank);
    }
    public <E> Set<E> reduceSetIntersectionProcess(Set<E> setData, int rootRank) throws Mp4jException {
        return process.reduceSetIntersection(setData, rootRank);
    }
    public <E> Map<String, List<E>> reduceMapListConcatProcess(Map<String, List<E>> mapData, int rootRank) throws Mp4jException {
        return process.reduceMapListConcat(mapData, rootRank);
    }
    public <E> List<E> reduceListConcatProcess(List<E> listData, int rootRank) throws Mp4jException {
        return process.reduceListConcat(listData, rootRank);
    }
    public <T> T allreduceArrayProcess(T arrData, Operand operand, IOperator operator, int from, int to) throws Mp4jException {
        return process.allreduceArray(arrData, operand, operator, from, to);
    }
    public <T> T allreduceArrayRpcProcess(T arrData, Operand operand, IOperator operator) throws Mp4jException {
        return process.allreduceArrayRpc(arrData, operand, operator);
    }
    public <T> T allreduceProcess(T value, Operand operand, IOperator operator) throws Mp4jException {
        return process.allreduce(value, operand, operator);
    }
    public <T> T allreduceRpcProcess(T value, Operand operand, IOperator operator) throws Mp4jException {
        return process.allreduceRpc(value, operand, operator);
    }
    public <V> Map<String, V> allreduceMapProcess(Map<String, V> mapData, Operand operand, IOperator operator) throws Mp4jException {
        return process.allreduceMap(mapData, operand, operator);
    }
    public <E> Map<String, Set<E>> allreduceMapSetUnionProcess(Map<String, Set<E>> mapData) throws Mp4jException {
        return process.allreduceMapSetUnion(mapData);
    }
    public <E> Set<E> allreduceSetUnionProcess(Set<E> setData) throws Mp4jException {
        return process.allreduceSetUnion(setData);
    }
    public <E> Map<String, Set<E>> allreduceMapSetIntersectionProcess(Map<String, Set<E>> mapData) throws Mp4jException {
        return process.allreduceMapSetIntersection(mapData);
    }
    public <E> Set<E> allreduceSetIntersectionProcess(Set<E> setData) throws Mp4jException {
        return process.allreduceSetIntersection(setData);
    }
    public <E> Map<String, List<E>> allreduceMapListConcatProcess(Map<String, List<E>> mapData) throws Mp4jException {
        return process.allreduceMapListConcat(mapData);
    }
    public <E> List<E> allreduceListConcatProcess(List<E> listData) throws Mp4jException {
        return process.allreduceListConcat(listData);
    }
    // ------------------------------------------------------------------ device buffers (GPU plane)
    // The thread-level collectives on raw CUDA addresses (one buffer per thread, e.g. from a CUDA binding of the
    // caller's choice). Every thread calls with its own buffer; deposit, launch by one thread and delivery of its
    // error to all threads are native (mp4j_thread_gpu_*, mp4j::ThreadComm::delegate). One fused kernel per call:
    // thread reduce + inter-GPU step + fan-out (allreduce: csrc/gpu/thread_fused_ws.cu, the others: thread_seg.cu).
    private volatile long tgroup = 0;

    /** call once (any one thread) before the ...Device methods */
    public synchronized void initGpu(int device, long heapBytes, long stagingBytes) throws Mp4jException {
        process.initGpu(device, heapBytes, stagingBytes);
        if (tgroup == 0) tgroup = Native.tgroupCreate(threadNum);
        if (tgroup == 0) throw new Mp4jException(Native.lastError());
    }

    private long plane() throws Mp4jException {
        if (tgroup == 0 || process.gpuHandle() == 0) throw new Mp4jException("call initGpu first");
        return process.gpuHandle();
    }

    private void segments(String name, long devPtr, Operand operand, int op, long[] segs, int launchTid, long stream) throws Mp4jException {
        if (Native.threadGpuSegments(tgroup, plane(), getThreadId(), launchTid, devPtr, operand.dtype, op, segs, name, stream) != 0)
            throw new Mp4jException(Native.lastError());
    }

    /** segment table rows: lo, hi, srcRank, srcThread, dstRank, dstThread (-1 = all) */
    private long[] table(int[][] lo, int[][] hi, boolean fromAll, int srcRank, int srcThread, int dstRank, int dstThread) {
        int p = getSlaveNum();
        long[] segs = new long[p * threadNum * 6];
        int k = 0;
        for (int q = 0; q < p; q++) for (int t = 0; t < threadNum; t++) {
            segs[k++] = lo[q][t];
            segs[k++] = hi[q][t];
            segs[k++] = fromAll ? -1 : (srcRank == -2 ? q : srcRank);
            segs[k++] = fromAll ? -1 : (srcRank == -2 ? t : srcThread);
            segs[k++] = dstRank == -2 ? q : dstRank;
            segs[k++] = dstRank == -2 ? t : dstThread;
        }
        return segs;
    }

    private int launcher(int rootRank, int rootThreadId) { return getRank() == rootRank ? rootThreadId : 0; }  // TCS:486

    public void allreduceArrayDevice(long devPtr, Operand operand, IOperator operator, long from, long to, long stream) throws Mp4jException {
        if (from < 0 || to < from) throw new Mp4jException("from must be >= 0 and <= to!");
        if (Native.threadGpuAllreduce(tgroup, plane(), getThreadId(), devPtr, operand.dtype, ProcessCommSlave.deviceOp(operator), from, to, stream) != 0)
            throw new Mp4jException(Native.lastError());
    }

    /** thread (q, t) ends with the reduced [from + counts before it, + counts[q][t]) (TCS:1174) */
    public void reduceScatterArrayDevice(long devPtr, Operand operand, IOperator operator, int from, int[][] counts, long stream) throws Mp4jException {
        CommUtils.isFromCountsLegal(from, counts);
        int p = getSlaveNum();
        int[][] lo = new int[p][threadNum], hi = new int[p][threadNum];
        int off = from;
        for (int q = 0; q < p; q++) for (int t = 0; t < threadNum; t++) {
            lo[q][t] = off;
            off += counts[q][t];
            hi[q][t] = off;
        }
        segments("thread_reduce_scatter", devPtr, operand, ProcessCommSlave.deviceOp(operator), table(lo, hi, true, 0, 0, -2, 0), 0, stream);
    }

    public void reduceArrayDevice(long devPtr, Operand operand, IOperator operator, int from, int to, int rootRank, int rootThreadId, long stream) throws Mp4jException {
        CommUtils.isFromToLegal(from, to);
        segments("thread_reduce", devPtr, operand, ProcessCommSlave.deviceOp(operator), new long[]{from, to, -1, -1, rootRank, rootThreadId},
                launcher(rootRank, rootThreadId), stream);
    }

    public void broadcastArrayDevice(long devPtr, Operand operand, int from, int to, int rootRank, int rootThreadId, long stream) throws Mp4jException {
        CommUtils.isFromToLegal(from, to);
        segments("thread_broadcast", devPtr, operand, 0 /* MP4J_COPY is ignored: no segment reduces */, new long[]{from, to, rootRank, rootThreadId, -1, -1},
                launcher(rootRank, rootThreadId), stream);
    }

    public void allgatherArrayDevice(long devPtr, Operand operand, int[][] froms, int[][] tos, long stream) throws Mp4jException {
        CommUtils.isfromsTosLegal(froms, tos, threadNum);
        segments("thread_allgather", devPtr, operand, 0, table(froms, tos, false, -2, 0, -1, -1), 0, stream);
    }

    public void gatherArrayDevice(long devPtr, Operand operand, int[][] sendfroms, int[][] sendtos, int rootRank, int rootThreadId, long stream) throws Mp4jException {
        CommUtils.isfromsTosLegal(sendfroms, sendtos, threadNum);
        segments("thread_gather", devPtr, operand, 0, table(sendfroms, sendtos, false, -2, 0, rootRank, rootThreadId),
                launcher(rootRank, rootThreadId), stream);
    }

    public void scatterArrayDevice(long devPtr, Operand operand, int[][] recvfroms, int[][] recvtos, int rootRank, int rootThreadId, long stream) throws Mp4jException {
        CommUtils.isfromsTosLegal(recvfroms, recvtos, threadNum);
        segments("thread_scatter", devPtr, operand, 0, table(recvfroms, recvtos, false, rootRank, rootThreadId, -2, 0),
                launcher(rootRank, rootThreadId), stream);
    }
}
