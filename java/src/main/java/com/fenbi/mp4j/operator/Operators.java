package com.fenbi.mp4j.operator;

/** Predefined operators, grouped like the reference's Operators class (Double / Float / Long / Int /
 *  Short / Byte). Every constant implements the typed interface of its group - so it can be applied
 *  element by element on the Java side (thread-level combine, maps) - and carries the native op code
 *  (mp4j_op of include/mp4j_types.h) under which the C++ host collectives and the CUDA kernels run the
 *  same reduction (csrc/core/ops.h).
 *
 *  Semantics: SUM / PROD of short and byte wrap on narrowing; floating MAX / MIN are
 *  Math.max / Math.min (NaN-propagating, -0.0 &lt; +0.0). Deviation from the reference, shared with the
 *  native core: the packed LOC operators break ties by the LOWER location (the reference keeps the
 *  first argument), which makes them commutative so that every rank computes identical bits. */
public final class Operators {
    private Operators() {}

    static final int SUM = 0, MAX = 1, MIN = 2, PROD = 3, BAND = 4, BOR = 5, BXOR = 6, MAXLOC = 7, MINLOC = 8;

    private abstract static class Base implements IOperator {
        final int code;
        final String name;
        Base(int code, String name) { this.code = code; this.name = name; }
        @Override public int nativeCode() { return code; }
        @Override public String toString() { return name; }
    }

    public static final class Double {
        private Double() {}
        private static final class Op extends Base implements IDoubleOperator {
            Op(int code, String name) { super(code, name); }
            @Override public double apply(double a, double b) {
                switch (code) {
                    case Operators.SUM: return a + b;
                    case Operators.MAX: return Math.max(a, b);
                    case Operators.MIN: return Math.min(a, b);
                    case Operators.PROD: return a * b;
                    case Operators.MAXLOC: return loc(a, b, true);
                    default: return loc(a, b, false);
                }
            }
            private static double loc(double a, double b, boolean max) {
                float va = getFloatVal(a), vb = getFloatVal(b);
                if (max ? va > vb : va < vb) return a;
                if (max ? vb > va : vb < va) return b;
                if (va != va) return a;
                if (vb != vb) return b;
                return getIntLoc(a) <= getIntLoc(b) ? a : b;
            }
        }
        public static final IDoubleOperator SUM = new Op(Operators.SUM, "SUM");
        public static final IDoubleOperator MAX = new Op(Operators.MAX, "MAX");
        public static final IDoubleOperator MIN = new Op(Operators.MIN, "MIN");
        public static final IDoubleOperator PROD = new Op(Operators.PROD, "PROD");
        /** value = float in the high 32 bits, location = int in the low 32 bits of the double */
        public static final IDoubleOperator FLOAT_MAX_LOC = new Op(MAXLOC, "FLOAT_MAX_LOC");
        public static final IDoubleOperator FLOAT_MIN_LOC = new Op(MINLOC, "FLOAT_MIN_LOC");

        public static double compositeDouble(float val, int loc) {
            long ret = (0xffffffffL & java.lang.Float.floatToRawIntBits(val)) << 32 | (0xffffffffL & loc);
            return java.lang.Double.longBitsToDouble(ret);
        }
        public static float getFloatVal(double composite) {
            return java.lang.Float.intBitsToFloat((int) (java.lang.Double.doubleToRawLongBits(composite) >>> 32));
        }
        public static int getIntLoc(double composite) {
            return (int) (java.lang.Double.doubleToRawLongBits(composite) & 0xffffffffL);
        }
    }

    public static final class Float {
        private Float() {}
        private static final class Op extends Base implements IFloatOperator {
            Op(int code, String name) { super(code, name); }
            @Override public float apply(float a, float b) {
                switch (code) {
                    case Operators.SUM: return a + b;
                    case Operators.MAX: return Math.max(a, b);
                    case Operators.MIN: return Math.min(a, b);
                    default: return a * b;
                }
            }
        }
        public static final IFloatOperator SUM = new Op(Operators.SUM, "SUM");
        public static final IFloatOperator MAX = new Op(Operators.MAX, "MAX");
        public static final IFloatOperator MIN = new Op(Operators.MIN, "MIN");
        public static final IFloatOperator PROD = new Op(Operators.PROD, "PROD");
    }

    public static final class Long {
        private Long() {}
        private static final class Op extends Base implements ILongOperator {
            Op(int code, String name) { super(code, name); }
            @Override public long apply(long a, long b) {
                switch (code) {
                    case Operators.SUM: return a + b;
                    case Operators.MAX: return Math.max(a, b);
                    case Operators.MIN: return Math.min(a, b);
                    case Operators.PROD: return a * b;
                    case Operators.BAND: return a & b;
                    case Operators.BOR: return a | b;
                    case Operators.BXOR: return a ^ b;
                    case Operators.MAXLOC: return loc(a, b, true);
                    default: return loc(a, b, false);
                }
            }
            private static long loc(long a, long b, boolean max) {
                int va = getIntVal(a), vb = getIntVal(b);
                if (va != vb) return (max ? va > vb : va < vb) ? a : b;
                return getIntLoc(a) <= getIntLoc(b) ? a : b;
            }
        }
        public static final ILongOperator SUM = new Op(Operators.SUM, "SUM");
        public static final ILongOperator MAX = new Op(Operators.MAX, "MAX");
        public static final ILongOperator MIN = new Op(Operators.MIN, "MIN");
        public static final ILongOperator PROD = new Op(Operators.PROD, "PROD");
        public static final ILongOperator BITS_AND = new Op(BAND, "BITS_AND");
        public static final ILongOperator BITS_OR = new Op(BOR, "BITS_OR");
        public static final ILongOperator BITS_XOR = new Op(BXOR, "BITS_XOR");
        /** value = int in the high 32 bits, location = int in the low 32 bits of the long */
        public static final ILongOperator INT_MAX_LOC = new Op(MAXLOC, "INT_MAX_LOC");
        public static final ILongOperator INT_MIN_LOC = new Op(MINLOC, "INT_MIN_LOC");

        public static long compositeLong(int val, int loc) { return ((long) val << 32) | (0xffffffffL & loc); }
        public static int getIntVal(long composite) { return (int) (composite >>> 32); }
        public static int getIntLoc(long composite) { return (int) (composite & 0xffffffffL); }
    }

    public static final class Int {
        private Int() {}
        private static final class Op extends Base implements IIntOperator {
            Op(int code, String name) { super(code, name); }
            @Override public int apply(int a, int b) {
                switch (code) {
                    case Operators.SUM: return a + b;
                    case Operators.MAX: return Math.max(a, b);
                    case Operators.MIN: return Math.min(a, b);
                    case Operators.PROD: return a * b;
                    case Operators.BAND: return a & b;
                    case Operators.BOR: return a | b;
                    default: return a ^ b;
                }
            }
        }
        public static final IIntOperator SUM = new Op(Operators.SUM, "SUM");
        public static final IIntOperator MAX = new Op(Operators.MAX, "MAX");
        public static final IIntOperator MIN = new Op(Operators.MIN, "MIN");
        public static final IIntOperator PROD = new Op(Operators.PROD, "PROD");
        public static final IIntOperator BITS_AND = new Op(BAND, "BITS_AND");
        public static final IIntOperator BITS_OR = new Op(BOR, "BITS_OR");
        public static final IIntOperator BITS_XOR = new Op(BXOR, "BITS_XOR");
    }

    public static final class Short {
        private Short() {}
        private static final class Op extends Base implements IShortOperator {
            Op(int code, String name) { super(code, name); }
            @Override public short apply(short a, short b) {
                switch (code) {
                    case Operators.SUM: return (short) (a + b);
                    case Operators.MAX: return (short) Math.max(a, b);
                    case Operators.MIN: return (short) Math.min(a, b);
                    case Operators.PROD: return (short) (a * b);
                    case Operators.BAND: return (short) (a & b);
                    case Operators.BOR: return (short) (a | b);
                    default: return (short) (a ^ b);
                }
            }
        }
        public static final IShortOperator SUM = new Op(Operators.SUM, "SUM");
        public static final IShortOperator MAX = new Op(Operators.MAX, "MAX");
        public static final IShortOperator MIN = new Op(Operators.MIN, "MIN");
        public static final IShortOperator PROD = new Op(Operators.PROD, "PROD");
        public static final IShortOperator BITS_AND = new Op(BAND, "BITS_AND");
        public static final IShortOperator BITS_OR = new Op(BOR, "BITS_OR");
        public static final IShortOperator BITS_XOR = new Op(BXOR, "BITS_XOR");
    }

    public static final class Byte {
        private Byte() {}
        private static final class Op extends Base implements IByteOperator {
            Op(int code, String name) { super(code, name); }
            @Override public byte apply(byte a, byte b) {
                switch (code) {
                    case Operators.SUM: return (byte) (a + b);
                    case Operators.MAX: return (byte) Math.max(a, b);
                    case Operators.MIN: return (byte) Math.min(a, b);
                    case Operators.PROD: return (byte) (a * b);
                    case Operators.BAND: return (byte) (a & b);
                    case Operators.BOR: return (byte) (a | b);
                    default: return (byte) (a ^ b);
                }
            }
        }
        public static final IByteOperator SUM = new Op(Operators.SUM, "SUM");
        public static final IByteOperator MAX = new Op(Operators.MAX, "MAX");
        public static final IByteOperator MIN = new Op(Operators.MIN, "MIN");
        public static final IByteOperator PROD = new Op(Operators.PROD, "PROD");
        public static final IByteOperator BITS_AND = new Op(BAND, "BITS_AND");
        public static final IByteOperator BITS_OR = new Op(BOR, "BITS_OR");
        public static final IByteOperator BITS_XOR = new Op(BXOR, "BITS_XOR");
    }
}
