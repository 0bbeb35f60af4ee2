"""B200-native KMeans (Lloyd fit / predict) behind the spark-rapids-ml plugin surface.

Sub-modules: ``_native`` (ctypes binding of the C ABI in include/b200km.h), ``build`` (nvcc build of
libb200km.so for sm_100a).  The PySpark-compatible classes live in ``clustering``.
"""
__version__ = "0.1.0"
