/* TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
 * Stand-in for <starpu.h> so that the reference's own src/core/tensor_builder.cpp compiles from where it lies
 * under /root/reference (StarPU 1.4.8 is not installed in this image).  It declares the one type that
 * translation unit uses, with the field order of StarPU 1.4's starpu_variable_interface
 * (include/starpu_data_interfaces.h); the reference's tests hand-declare the same struct around raw buffers
 * (/root/reference/tests/common/test_utils.hpp:88-135). */
#ifndef ORACLE_REF_STUB_STARPU_H
#define ORACLE_REF_STUB_STARPU_H
#include <stddef.h>
#include <stdint.h>
enum starpu_data_interface_id { STARPU_UNKNOWN_INTERFACE_ID = -1, STARPU_MATRIX_INTERFACE_ID = 0, STARPU_BLOCK_INTERFACE_ID = 1,
                                STARPU_VECTOR_INTERFACE_ID = 2, STARPU_CSR_INTERFACE_ID = 3, STARPU_BCSR_INTERFACE_ID = 4,
                                STARPU_VARIABLE_INTERFACE_ID = 5 };
struct starpu_variable_interface {
  enum starpu_data_interface_id id;
  uintptr_t ptr;
  uintptr_t dev_handle;
  size_t offset;
  size_t elemsize;
};
#endif
