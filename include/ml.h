/*
 * ml.h -- alias of model_runner.h for callers that include the reference's
 * older header name (model_runner/ml.h).
 *
 * The reference's ml.h differs from its model_runner.h only by two extra
 * trailing fields in ml_model_params (gpu_memory_fraction ml.h:61,
 * visible_devices ml.h:67) and by lacking ML_FLOAT16 (ml.h:103-106).  The
 * struct is a strict prefix extension and this library reads only the first
 * three fields (as the reference does: model.cpp:80,101,107), so a caller
 * compiled against either layout is ABI compatible (sizeof 24 vs 40, LP64).
 * Device selection never comes from this struct: use the environment
 * (CUDA_VISIBLE_DEVICES / MLB200_DEVICE), see INTEGRATION.md.
 */
#ifndef ML_H_B200
#define ML_H_B200
#include "model_runner.h"
#endif /* ML_H_B200 */
