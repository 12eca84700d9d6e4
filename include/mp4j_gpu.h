/* GPU half of the C ABI (filled in by csrc/gpu). Included from mp4j.h inside extern "C". */
#ifndef MP4J_GPU_H_
#define MP4J_GPU_H_
#endif
