// Library-wide C-ABI entry points: version and last-error string.
#include "tdl_common.cuh"

TDL_API const char* tdl_last_error() { return tdl::err_buf(); }

// Encodes the ABI revision; bump when an existing signature changes.
TDL_API int tdl_version() { return 100; }

// Number of SMs of the current device as the kernels' launch heuristics see it.
TDL_API int tdl_num_sms() { return tdl::num_sms(); }
