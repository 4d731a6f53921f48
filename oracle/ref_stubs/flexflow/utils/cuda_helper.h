// Stand-in for the reference header of the same name when its kernels are compiled into oracle/_ref/ (see
// oracle/ref_decompress_wrap.cu, which supplies the three names the kernels use).  Intentionally empty.
