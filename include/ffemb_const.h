/* ffemb_const.h -- enum values of the reference, unchanged.
 *
 * Same names and numeric values as the reference's include/ffconst.h:11-28
 * (AggrMode) and :22-28 (DataType), so that a binding written against the
 * reference's constants keeps working.  Plain C; parseable by `gcc -E | cffi`.
 */
#ifndef FFEMB_CONST_H
#define FFEMB_CONST_H

enum AggrMode {
  AGGR_MODE_NONE = 20,
  AGGR_MODE_SUM = 21,
  AGGR_MODE_AVG = 22,
};

enum DataType {
  DT_FLOAT = 40,
  DT_DOUBLE = 41,
  DT_INT32 = 42,
  DT_INT64 = 43,
  DT_BOOLEAN = 44,
};

/* FFProtoBuf.Op.DeviceType / MemoryType, src/runtime/strategy.proto:7-14 */
enum FFEmbDeviceType {
  FFEMB_DEVICE_GPU = 0,
  FFEMB_DEVICE_CPU = 1,
};

enum FFEmbMemoryType {
  FFEMB_MEMORY_FBM = 0,
  FFEMB_MEMORY_ZCM = 1,
};

/* Status codes returned by every ffemb_* call that can fail (the reference
 * asserts / exit(1)s instead: include/cuda_helper.h:6-36). */
enum FFEmbStatus {
  FFEMB_OK = 0,
  FFEMB_ERR_INVALID_ARGUMENT = 1,
  FFEMB_ERR_CUDA = 2,
  FFEMB_ERR_INDEX_OUT_OF_RANGE = 3,
  FFEMB_ERR_UNSUPPORTED = 4, /* e.g. momentum / weight decay != 0 */
  FFEMB_ERR_WORKSPACE_TOO_SMALL = 5,
  FFEMB_ERR_IO = 6,
  FFEMB_ERR_COMM = 7,
};

#endif /* FFEMB_CONST_H */
