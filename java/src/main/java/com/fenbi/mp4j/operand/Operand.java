package com.fenbi.mp4j.operand;

import com.fenbi.mp4j.exception.Mp4jException;
import com.fenbi.mp4j.operator.Collective;
import com.fenbi.mp4j.operator.Container;
import com.fenbi.mp4j.operator.IOperator;
import com.fenbi.mp4j.utils.ArrayOps;

import java.util.Map;

/** Typed payload descriptor (parity: J/operand/Operand.java:42-80 and the per-type subclasses).
 *
 *  In the reference an operand is also the wire codec (four Kryo serializers per type) and carries
 *  the per-call state (collective, container, operator). Here the wire work is native: `dtype` selects
 *  the templated C++/CUDA code path for the six primitive types, and String/Object payloads travel as
 *  serialized bytes over the native byte transport. The per-call state accessors are kept for source
 *  compatibility; the communicators do not rely on them, so one operand may be shared between calls.
 *  The thread-level helpers work on plain arrays / maps instead of the reference's MetaData wrappers. */
public abstract class Operand {
    /** mp4j_dtype of include/mp4j_types.h, or -1 for String/Object payloads */
    public final int dtype;
    public final String name;
    protected boolean compress;
    protected Collective collective;
    protected Container container;
    protected IOperator operator;

    protected Operand(String name, int dtype, boolean compress) {
        this.name = name;
        this.dtype = dtype;
        this.compress = compress;
    }

    public boolean isPrimitive() { return dtype >= 0; }
    public boolean isCompress() { return compress; }
    public void setCompress(boolean compress) { this.compress = compress; }
    public Collective getCollective() { return collective; }
    public void setCollective(Collective collective) { this.collective = collective; }
    public Container getContainer() { return container; }
    public void setContainer(Container container) { this.container = container; }
    public IOperator getOperator() { return operator; }
    public void setOperator(IOperator operator) { this.operator = operator; }

    /** A new array of this operand's element type. */
    public abstract Object newArray(int length);

    // ---- thread-level helpers (DOP:362-482 and siblings): arrays of this operand's element type
    /** to[start, end) = from[start, end) */
    public void threadCopy(Object from, Object to, int start, int end) { ArrayOps.copy(from, to, start, end); }
    /** to = from over the whole array (non-delegate threads take the full result, TCS:676) */
    public void threadArrayAllCopy(Object from, Object to) { ArrayOps.copy(from, to, 0, ArrayOps.length(from)); }
    /** to[i] = op(to[i], from[i]) for i in [start, end) (DOP:441-457) */
    public void threadReduce(Object from, Object to, int start, int end, IOperator op) throws Mp4jException {
        ArrayOps.reduce(to, from, start, end, 0, op);
    }
    /** to[k] = op(to[k], from[k]) or from[k] if absent (DOP:458-477) */
    public <V> void threadMerge(Map<String, V> from, Map<String, V> to, IOperator op) throws Mp4jException {
        ArrayOps.mergeMap(to, from, op);
    }

    /** Bytes of a payload of this operand (array slice, map, list of maps ...) for the byte transport. */
    public byte[] toBytes(Object payload) throws Mp4jException { return ArrayOps.serialize(payload, compress); }
    public <X> X fromBytes(byte[] bytes) throws Mp4jException { return ArrayOps.deserialize(bytes, compress); }

    /** Stream forms of the codec (the reference's send / recv write Kryo streams bound to its MetaData
     *  headers; here a payload is one length-prefixed block of {@link #toBytes} output). */
    public void send(java.io.OutputStream out, Object payload) throws Mp4jException {
        try {
            byte[] b = toBytes(payload);
            java.io.DataOutputStream d = new java.io.DataOutputStream(out);
            d.writeInt(b.length);
            d.write(b);
            d.flush();
        } catch (java.io.IOException e) {
            throw new Mp4jException(e);
        }
    }
    public <X> X recv(java.io.InputStream in) throws Mp4jException {
        try {
            java.io.DataInputStream d = new java.io.DataInputStream(in);
            byte[] b = new byte[d.readInt()];
            d.readFully(b);
            return this.<X>fromBytes(b);
        } catch (java.io.IOException e) {
            throw new Mp4jException(e);
        }
    }

    @Override public String toString() { return name + (compress ? "(compress)" : ""); }
}
