package com.fenbi.mp4j.comm;

import com.fenbi.mp4j.exception.Mp4jException;
import com.fenbi.mp4j.operand.Operand;
import com.fenbi.mp4j.operator.IOperator;

import java.util.concurrent.CyclicBarrier;

/** Process x thread communicator: threads deposit their arrays, one delegate reduces them and runs
 *  the process collective, the others copy the result out (same delegate pattern as the Python
 *  binding; the GPU form of this class is the fused kernel of csrc/gpu/thread_fused_ws.cu). */
public class ThreadCommSlave {
    private final ProcessCommSlave process;
    private final int threadNum;
    private final CyclicBarrier barrier;
    private final ThreadLocal<Integer> threadId = ThreadLocal.withInitial(() -> 0);
    private final Object[] slots;

    public ThreadCommSlave(String loginName, int threadNum, String masterHost, int masterPort) throws Mp4jException {
        this.process = new ProcessCommSlave(loginName, masterHost, masterPort);
        this.threadNum = threadNum;
        this.barrier = new CyclicBarrier(threadNum);
        this.slots = new Object[threadNum];
    }
    public int getThreadNum() { return threadNum; }
    public void setThreadId(int id) { threadId.set(id); }
    public int getThreadId() { return threadId.get(); }
    public int getRank() { return process.getRank(); }
    public int getSlaveNum() { return process.getSlaveNum(); }
    public void threadBarrier() throws Mp4jException {
        try { barrier.await(); } catch (Exception e) { throw new Mp4jException(e); }
    }
    public void barrier() throws Mp4jException {
        threadBarrier();
        if (getThreadId() == 0) process.barrier();
        threadBarrier();
    }
    public void info(String info) throws Mp4jException { if (getThreadId() == 0) process.info(info); }
    public void close(int code) throws Mp4jException { process.close(code); }

    /** pass-through for "one thread per process calls this" */
    public double[] allreduceArrayProcess(double[] arr, Operand operand, IOperator operator, int from, int to) throws Mp4jException {
        return process.allreduceArray(arr, operand, operator, from, to);
    }

    /** all slaveNum x threadNum arrays are reduced; every thread's array holds the result */
    public double[] allreduceArray(double[] arr, Operand operand, IOperator operator, int from, int to) throws Mp4jException {
        if (threadNum == 1) return process.allreduceArray(arr, operand, operator, from, to);
        int tid = getThreadId();
        slots[tid] = arr;
        threadBarrier();
        if (tid == 0) {
            double[] acc = (double[]) slots[0];
            int code = operator.nativeCode();
            for (int t = 1; t < threadNum; t++) {
                double[] src = (double[]) slots[t];
                for (int i = from; i < to; i++) {
                    switch (code) {
                        case 0: acc[i] += src[i]; break;
                        case 1: acc[i] = Math.max(acc[i], src[i]); break;
                        case 2: acc[i] = Math.min(acc[i], src[i]); break;
                        case 3: acc[i] *= src[i]; break;
                        default: throw new Mp4jException("operator not supported at thread level in the Java facade");
                    }
                }
            }
            process.allreduceArray(acc, operand, operator, from, to);
        }
        threadBarrier();
        if (tid != 0) System.arraycopy((double[]) slots[0], from, arr, from, to - from);
        threadBarrier();
        return arr;
    }
}
