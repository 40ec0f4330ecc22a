// dtype.h -- element types of ml_image buffers (reference dtype.h:40-53; the
// TensorFlow enum mapping of dtype.h:10-38 becomes the GraphDef DataType
// numbers DT_FLOAT=1 / DT_HALF=19, handled in graphdef.cpp).
#pragma once

#include "model_runner.h"

#include <stdexcept>
#include <string>

namespace ML {

inline size_t DataTypeSize(ml_data_type type)
{
    if (type == ML_FLOAT32) return 4;
    if (type == ML_FLOAT16) return 2;
    throw std::runtime_error("Unsupported image data type: " + std::to_string(static_cast<int>(type)));
}

// GraphDef DataType enum values (tensorflow/core/framework/types.proto).
constexpr int kTfFloat = 1;
constexpr int kTfHalf = 19;

inline ml_data_type DataTypeFromTF(int tf_dtype)
{
    if (tf_dtype == kTfFloat) return ML_FLOAT32;
    if (tf_dtype == kTfHalf) return ML_FLOAT16;
    throw std::runtime_error("Unsupported image data type: " + std::to_string(tf_dtype));
}

inline int DataTypeToTF(ml_data_type type)
{
    if (type == ML_FLOAT32) return kTfFloat;
    if (type == ML_FLOAT16) return kTfHalf;
    throw std::runtime_error("Unsupported image data type: " + std::to_string(static_cast<int>(type)));
}

} // namespace ML
