package com.fenbi.mp4j.operator;

/** Predefined operators, grouped like the reference's Operators class. Each constant is a native
 *  op code: the reduction itself runs in libmp4j (csrc/core/ops.h), on the host or on the GPU. */
public final class Operators {
    private Operators() {}

    public static final class Native implements IOperator {
        private final int code;
        private final String name;
        Native(int code, String name) { this.code = code; this.name = name; }
        @Override public int nativeCode() { return code; }
        @Override public String toString() { return name; }
    }

    private static final int SUM = 0, MAX = 1, MIN = 2, PROD = 3, BAND = 4, BOR = 5, BXOR = 6, MAXLOC = 7, MINLOC = 8;

    public static final class Double {
        public static final IOperator SUM = new Native(Operators.SUM, "SUM");
        public static final IOperator MAX = new Native(Operators.MAX, "MAX");
        public static final IOperator MIN = new Native(Operators.MIN, "MIN");
        public static final IOperator PROD = new Native(Operators.PROD, "PROD");
        public static final IOperator FLOAT_MAX_LOC = new Native(MAXLOC, "FLOAT_MAX_LOC");
        public static final IOperator FLOAT_MIN_LOC = new Native(MINLOC, "FLOAT_MIN_LOC");

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
        public static final IOperator SUM = new Native(Operators.SUM, "SUM");
        public static final IOperator MAX = new Native(Operators.MAX, "MAX");
        public static final IOperator MIN = new Native(Operators.MIN, "MIN");
        public static final IOperator PROD = new Native(Operators.PROD, "PROD");
    }

    public static final class Long {
        public static final IOperator SUM = new Native(Operators.SUM, "SUM");
        public static final IOperator MAX = new Native(Operators.MAX, "MAX");
        public static final IOperator MIN = new Native(Operators.MIN, "MIN");
        public static final IOperator PROD = new Native(Operators.PROD, "PROD");
        public static final IOperator BITS_AND = new Native(BAND, "BITS_AND");
        public static final IOperator BITS_OR = new Native(BOR, "BITS_OR");
        public static final IOperator BITS_XOR = new Native(BXOR, "BITS_XOR");
        public static final IOperator INT_MAX_LOC = new Native(MAXLOC, "INT_MAX_LOC");
        public static final IOperator INT_MIN_LOC = new Native(MINLOC, "INT_MIN_LOC");

        public static long compositeLong(int val, int loc) { return ((long) val << 32) | (0xffffffffL & loc); }
        public static int getIntVal(long composite) { return (int) (composite >>> 32); }
        public static int getIntLoc(long composite) { return (int) (composite & 0xffffffffL); }
    }

    public static final class Int {
        public static final IOperator SUM = new Native(Operators.SUM, "SUM");
        public static final IOperator MAX = new Native(Operators.MAX, "MAX");
        public static final IOperator MIN = new Native(Operators.MIN, "MIN");
        public static final IOperator PROD = new Native(Operators.PROD, "PROD");
        public static final IOperator BITS_AND = new Native(BAND, "BITS_AND");
        public static final IOperator BITS_OR = new Native(BOR, "BITS_OR");
        public static final IOperator BITS_XOR = new Native(BXOR, "BITS_XOR");
    }

    public static final class Short {
        public static final IOperator SUM = new Native(Operators.SUM, "SUM");
        public static final IOperator MAX = new Native(Operators.MAX, "MAX");
        public static final IOperator MIN = new Native(Operators.MIN, "MIN");
        public static final IOperator PROD = new Native(Operators.PROD, "PROD");
        public static final IOperator BITS_AND = new Native(BAND, "BITS_AND");
        public static final IOperator BITS_OR = new Native(BOR, "BITS_OR");
        public static final IOperator BITS_XOR = new Native(BXOR, "BITS_XOR");
    }

    public static final class Byte {
        public static final IOperator SUM = new Native(Operators.SUM, "SUM");
        public static final IOperator MAX = new Native(Operators.MAX, "MAX");
        public static final IOperator MIN = new Native(Operators.MIN, "MIN");
        public static final IOperator PROD = new Native(Operators.PROD, "PROD");
        public static final IOperator BITS_AND = new Native(BAND, "BITS_AND");
        public static final IOperator BITS_OR = new Native(BOR, "BITS_OR");
        public static final IOperator BITS_XOR = new Native(BXOR, "BITS_XOR");
    }
}
