/* TEST INFRASTRUCTURE.  The reference's CUDA lattice (crf/lattice/cuda_lattice/permutohedral.cu)
 * links against NVIDIA's 2009 `libcutil.a` (a prebuilt x86 archive vendored beside it) for two
 * command-line helpers that only its optional initCuda() entry point reaches through the
 * CUT_DEVICE_INIT macro (cutil.h:820-828).  The benchmark never calls initCuda(); these
 * definitions just satisfy the dynamic linker.  Nothing here is derived from cutil's sources. */
int cutCheckCmdLineFlag(const int argc, const char** argv, const char* flag_name) {
  (void)argc; (void)argv; (void)flag_name;
  return 0; /* CUTFalse */
}
int cutGetCmdLineArgumenti(const int argc, const char** argv, const char* arg_name, int* val) {
  (void)argc; (void)argv; (void)arg_name; (void)val;
  return 0; /* CUTFalse: keep device 0 */
}
