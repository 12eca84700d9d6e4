/* mp4j-b200 public enums shared by the C ABI, the C++ core and the CUDA kernels.
 *
 * Parity: operand types of J/operand/Operands.java:33-107, operators of
 * J/operator/Operators.java:30-351, collectives/containers of
 * J/operator/Collective.java:29-36 and J/operator/Container.java:29-31.
 */
#ifndef MP4J_TYPES_H_
#define MP4J_TYPES_H_

#ifdef __cplusplus
extern "C" {
#endif

/* Element types. The first six are the reference's primitive operands
 * (double, float, long, int, short, byte); BF16/F16 are additions for torch users. */
typedef enum mp4j_dtype {
  MP4J_F64 = 0,
  MP4J_F32 = 1,
  MP4J_I64 = 2,
  MP4J_I32 = 3,
  MP4J_I16 = 4,
  MP4J_I8 = 5,
  MP4J_BF16 = 6,
  MP4J_F16 = 7,
  MP4J_DTYPE_COUNT = 8
} mp4j_dtype;

/* Reduction operators. MAXLOC/MINLOC are the packed (value, location) operators:
 * on F64 the payload is (float value << 32 | int loc) (Operators.Double.FLOAT_MAX_LOC),
 * on I64 it is (int value << 32 | int loc) (Operators.Long.INT_MAX_LOC).
 * COPY is "no reduction" (gather/scatter/allgather/broadcast). */
typedef enum mp4j_op {
  MP4J_SUM = 0,
  MP4J_MAX = 1,
  MP4J_MIN = 2,
  MP4J_PROD = 3,
  MP4J_BAND = 4,
  MP4J_BOR = 5,
  MP4J_BXOR = 6,
  MP4J_MAXLOC = 7,
  MP4J_MINLOC = 8,
  MP4J_OP_COUNT = 9,
  MP4J_COPY = 15
} mp4j_op;

/* Custom device operators (mp4j_gpu_custom_op): run-time codes start here; the kernels see them as
 * the one template value below. */
#define MP4J_CUSTOM_OP_BASE 64
#define MP4J_CUSTOM_OP_TEMPLATE 63

typedef enum mp4j_collective {
  MP4J_GATHER = 0,
  MP4J_SCATTER = 1,
  MP4J_ALL_GATHER = 2,
  MP4J_BROADCAST = 3,
  MP4J_REDUCE_SCATTER = 4,
  MP4J_REDUCE = 5,
  MP4J_ALL_REDUCE = 6
} mp4j_collective;

typedef enum mp4j_container { MP4J_ARRAY = 0, MP4J_MAP = 1 } mp4j_container;

/* Wire compression for float payloads on the GPU path (replaces the reference's
 * Deflate stream wrapper, DOP:267-269). LOSSLESS is the default. */
typedef enum mp4j_wire {
  MP4J_WIRE_LOSSLESS = 0,
  MP4J_WIRE_BF16 = 1,
  MP4J_WIRE_FP8 = 2 /* e4m3 with one f32 scale per 128-element block */
} mp4j_wire;

/* Algorithm selector for the dense GPU collectives. AUTO consults the tuning table. */
typedef enum mp4j_algo {
  MP4J_ALGO_AUTO = 0,
  MP4J_ALGO_ONESHOT = 1,
  MP4J_ALGO_TWOSHOT = 2,
  MP4J_ALGO_NVLS = 3
} mp4j_algo;

#define MP4J_MAX_RANKS 8
#define MP4J_MAX_THREADS 64

/* Return codes of the C ABI: 0 ok, <0 error (text via mp4j_last_error()). */
#define MP4J_OK 0
#define MP4J_ERR -1

#ifdef __cplusplus
}
#endif
#endif /* MP4J_TYPES_H_ */
