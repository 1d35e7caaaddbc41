"""sparknet_b200: B200-native implementation of SparkNet's data-parallel hot path behind the
reference's libccaffe C ABI.  The CUDA/C++ product is csrc/ -> libccaffe.so; the Python here is
the host-side mirror of the reference's JVM glue (Net.scala, ProtoLoader.scala, the app loop)."""
__version__ = "0.1.0"
