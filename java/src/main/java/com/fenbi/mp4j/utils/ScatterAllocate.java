package com.fenbi.mp4j.utils;

import java.util.ArrayList;
import java.util.HashMap;
import java.util.List;
import java.util.Map;

/** Scatter-tree planners. Every plan maps a source rank to its sends in order; a send is
 *  {src, dst, rankFrom, rankTo}: "src hands the slices of ranks rankFrom..rankTo (inclusive) to dst".
 *
 *  allocate()        - binomial tree rooted at `root` on relative ranks: what the native host plane runs
 *                      (csrc/host/process_comm.cc); every non-root rank receives exactly once.
 *  allocateHalving() - the reference's plan (J/utils/ScatterAllocate.java:58-111): the root seeds rank 0
 *                      with the lower half and rank p/2 with the upper half, then each range halves.
 *                      For a root outside {0, p/2} the root's own slice leaves and comes back.
 *  The GPU path needs neither: every rank pulls its slice from the root in one kernel. */
public final class ScatterAllocate {
    private ScatterAllocate() {}

    public static Map<Integer, List<List<Integer>>> allocate(int p, int root) {
        Map<Integer, List<List<Integer>>> plan = new HashMap<Integer, List<List<Integer>>>();
        int top = 1;
        while (top < p) top <<= 1;
        top >>= 1;
        for (int v = 0; v < p; v++) {
            int m = v == 0 ? top : ((v & -v) >> 1);
            for (; m >= 1; m >>= 1) {
                if (v + m >= p) continue;
                int hi = Math.min(v + 2 * m, p) - 1;
                add(plan, (v + root) % p, (v + m + root) % p, v + m, hi);
            }
        }
        return plan;
    }

    public static Map<Integer, List<List<Integer>>> allocateHalving(int p, int root) {
        List<Integer> flat = new ArrayList<Integer>();
        int mid = p / 2;
        if (p > 1) {
            if (root != 0) addFlat(flat, root, 0, 0, mid - 1);
            if (root != mid) addFlat(flat, root, mid, mid, p - 1);
            split(0, mid - 1, flat);
            split(mid, p - 1, flat);
        }
        Map<Integer, List<List<Integer>>> plan = new HashMap<Integer, List<List<Integer>>>();
        for (int i = 0; i < flat.size(); i += 4) add(plan, flat.get(i), flat.get(i + 1), flat.get(i + 2), flat.get(i + 3));
        return plan;
    }

    /** Halving schedule of [start, end]: appends {start, upper, upper, end} and recurses into both halves. */
    public static void split(int start, int end, List<Integer> retList) {
        if (start >= end) return;
        int upper = start + (end - start + 1) / 2;
        addFlat(retList, start, upper, upper, end);
        split(start, upper - 1, retList);
        split(upper, end, retList);
    }

    /** rank -> number of messages it receives under `plan`. */
    public static Map<Integer, Integer> recvNum(Map<Integer, List<List<Integer>>> plan) {
        Map<Integer, Integer> out = new HashMap<Integer, Integer>();
        for (List<List<Integer>> sends : plan.values())
            for (List<Integer> s : sends) out.merge(s.get(1), 1, Integer::sum);
        return out;
    }

    private static void addFlat(List<Integer> flat, int src, int dst, int from, int to) {
        flat.add(src);
        flat.add(dst);
        flat.add(from);
        flat.add(to);
    }

    private static void add(Map<Integer, List<List<Integer>>> plan, int src, int dst, int from, int to) {
        List<Integer> send = new ArrayList<Integer>(4);
        addFlat(send, src, dst, from, to);
        plan.computeIfAbsent(src, k -> new ArrayList<List<Integer>>()).add(send);
    }

    /** Self-check over every root of a 13-rank job: nobody but the root may receive more than once. */
    public static void main(String[] args) {
        int p = 13;
        for (int root = 0; root < p; root++) {
            for (Map<Integer, List<List<Integer>>> plan : java.util.Arrays.asList(allocate(p, root), allocateHalving(p, root))) {
                for (Map.Entry<Integer, Integer> e : recvNum(plan).entrySet())
                    if (e.getValue() > 1 && e.getKey() != root) throw new IllegalStateException("rank " + e.getKey() + " receives " + e.getValue());
            }
        }
        System.out.println("ScatterAllocate: " + p + " roots ok");
    }
}
