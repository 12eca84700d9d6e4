package com.fenbi.mp4j.utils;

import com.fenbi.mp4j.exception.Mp4jException;

import java.util.Arrays;

/** Range validators and balanced partitioners with the reference's names (J/utils/CommUtils.java:34-222).
 *  Partition rule (SURVEY 2.7-1): avg = size / n, the first n-1 participants take avg elements, the last
 *  one takes the remainder. The same rules exist natively in csrc/core/partition.h. */
public final class CommUtils {
    private CommUtils() {}

    public static void isfromsTosLegal(int[] froms, int[] tos) throws Mp4jException {
        if (froms.length != tos.length) throw new Mp4jException("froms, tos's length must be equal!");
        for (int i = 0; i < froms.length; i++) {
            if (froms[i] < 0) throw new Mp4jException("froms[i] must be >= 0!");
            if (tos[i] < 0) throw new Mp4jException("tos[i] must be >= 0!");
            if (froms[i] > tos[i]) throw new Mp4jException("froms[i] must be <= tos[i]!");
            if (i >= 1 && froms[i] < tos[i - 1]) throw new Mp4jException("froms[i] must be >= to[i - 1]");
        }
    }

    /** 2-D form: [slaveNum][threadNum] ranges, ordered within every process and across processes. */
    public static void isfromsTosLegal(int[][] froms, int[][] tos, int threadNum) throws Mp4jException {
        if (froms.length != tos.length) throw new Mp4jException("froms, tos's length must be equal!");
        for (int r = 0; r < froms.length; r++) {
            if (froms[r].length != threadNum) throw new Mp4jException("froms arrays must be array[slaveNum][threadNum]");
            if (tos[r].length != threadNum) throw new Mp4jException("tos arrays must be array[slaveNum][threadNum]");
            isfromsTosLegal(froms[r], tos[r]);
        }
        isfromsTosLegal(getProcessFroms(froms), getProcessTos(tos));
    }

    public static void isFromCountsLegal(int from, int[] counts) throws Mp4jException {
        if (from < 0) throw new Mp4jException("from must be >= 0!");
        for (int c : counts) if (c < 0) throw new Mp4jException("counts[i] must be >= 0!");
    }

    public static void isFromCountsLegal(int from, int[][] counts) throws Mp4jException {
        if (from < 0) throw new Mp4jException("from must be >= 0!");
        for (int[] row : counts) for (int c : row) if (c < 0) throw new Mp4jException("counts[i][j] must be >= 0!");
    }

    public static void isFromToLegal(int from, int to) throws Mp4jException {
        if (from < 0) throw new Mp4jException("from must be >= 0!");
        if (to < 0) throw new Mp4jException("to must be >= 0!");
        if (from > to) throw new Mp4jException("from must be <= to!");
    }

    /** The ranges of a collective must lie inside the array that is handed to native code: the reference gets an
     *  ArrayIndexOutOfBoundsException from the JVM there, a native memcpy would read past the buffer.
     *  `hi` is the largest exclusive end index the call touches. */
    public static void isRangeInArray(Object arr, long hi) throws Mp4jException {
        if (arr == null || !arr.getClass().isArray()) return;
        int len = java.lang.reflect.Array.getLength(arr);
        if (hi > len) throw new Mp4jException("range end " + hi + " is beyond the array length " + len);
    }

    public static long maxEnd(int[] tos) {
        long m = 0;
        for (int t : tos) m = Math.max(m, t);
        return m;
    }

    public static long maxEnd(int[][] tos) {
        long m = 0;
        for (int[] row : tos) m = Math.max(m, maxEnd(row));
        return m;
    }

    public static long endOfCounts(int from, int[] counts) {
        long e = from;
        for (int c : counts) e += c;
        return e;
    }

    public static long endOfCounts(int from, int[][] counts) {
        long e = from;
        for (int[] row : counts) for (int c : row) e += c;
        return e;
    }

    public static int[] getFromsFromCount(int from, int[] counts, int slaveNum) {
        int[] froms = new int[slaveNum];
        for (int i = 0; i < slaveNum; i++) {
            froms[i] = from;
            from += counts[i];
        }
        return froms;
    }

    public static int[] getTosFromCount(int from, int[] counts, int slaveNum) {
        int[] tos = new int[slaveNum];
        for (int i = 0; i < slaveNum; i++) {
            from += counts[i];
            tos[i] = from;
        }
        return tos;
    }

    /** Collapses 2-D thread ranges to the range each process covers. */
    public static int[] getProcessFroms(int[][] froms) {
        int[] out = new int[froms.length];
        for (int r = 0; r < froms.length; r++) out[r] = froms[r][0];
        return out;
    }

    public static int[] getProcessTos(int[][] tos) {
        int[] out = new int[tos.length];
        for (int r = 0; r < tos.length; r++) out[r] = tos[r][tos[r].length - 1];
        return out;
    }

    public static int[] createProcessArrayFroms(int size, int slaveNum) {
        int avg = size / slaveNum;
        int[] froms = new int[slaveNum];
        for (int r = 0; r < slaveNum; r++) froms[r] = r * avg;
        return froms;
    }

    public static int[] createProcessArrayTos(int size, int slaveNum) {
        int avg = size / slaveNum;
        int[] tos = new int[slaveNum];
        for (int r = 0; r < slaveNum; r++) tos[r] = r == slaveNum - 1 ? size : (r + 1) * avg;
        return tos;
    }

    public static int[][] createThreadArrayFroms(int size, int slaveNum, int threadNum) {
        int[] pf = createProcessArrayFroms(size, slaveNum), pt = createProcessArrayTos(size, slaveNum);
        int[][] out = new int[slaveNum][];
        for (int r = 0; r < slaveNum; r++) {
            out[r] = createProcessArrayFroms(pt[r] - pf[r], threadNum);
            for (int t = 0; t < threadNum; t++) out[r][t] += pf[r];
        }
        return out;
    }

    public static int[][] createThreadArrayTos(int size, int slaveNum, int threadNum) {
        int[] pf = createProcessArrayFroms(size, slaveNum), pt = createProcessArrayTos(size, slaveNum);
        int[][] out = new int[slaveNum][];
        for (int r = 0; r < slaveNum; r++) {
            out[r] = createProcessArrayTos(pt[r] - pf[r], threadNum);
            for (int t = 0; t < threadNum; t++) out[r][t] += pf[r];
        }
        return out;
    }

    /** Prints a few partitions (the reference's CommUtils.main does the same kind of eyeball check). */
    public static void main(String[] args) throws Mp4jException {
        int[][] cases = {{10, 3, 2}, {1000, 8, 8}, {7, 8, 1}};
        for (int[] c : cases) {
            System.out.println("size=" + c[0] + " slaves=" + c[1] + " froms=" + Arrays.toString(createProcessArrayFroms(c[0], c[1]))
                    + " tos=" + Arrays.toString(createProcessArrayTos(c[0], c[1])));
            isfromsTosLegal(createThreadArrayFroms(c[0], c[1], c[2]), createThreadArrayTos(c[0], c[1], c[2]), c[2]);
        }
    }
}
