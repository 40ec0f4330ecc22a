"""simpleraytracer_b200 -- B200-native (sm_100a) implementation of the per-pixel
hot path of achernigin1987/SimpleRayTracer: the model_runner C API
(include/model_runner.h) executed by hand-written CUDA kernels, plus the CUDA
render loop behind include/srt_render.h.  The product is
simpleraytracer_b200/lib/libModelRunner.so (C ABI); this package is only its
ctypes binding and a GraphDef file writer for synthetic models.
"""
from . import api, graphdef_writer  # noqa: F401
from .api import (Context, Image, Model, Scene, MLError, load_library, library_path,  # noqa: F401
                  make_params, gen_cornell_box, gen_mesh)
