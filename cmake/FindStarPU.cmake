# Locate StarPU 1.4 (headers under <prefix>/include/starpu/1.4, libstarpu-1.4).  Hints: STARPU_DIR (cache or
# environment), then pkg-config `starpu-1.4`.  Defines the imported target StarPU::starpu.
include_guard(GLOBAL)
if(NOT STARPU_DIR AND DEFINED ENV{STARPU_DIR})
  set(STARPU_DIR $ENV{STARPU_DIR} CACHE PATH "StarPU installation prefix")
endif()
find_package(PkgConfig QUIET)
if(PkgConfig_FOUND AND NOT STARPU_DIR)
  pkg_check_modules(PC_STARPU QUIET starpu-1.4)
endif()
find_path(STARPU_INCLUDE_DIR starpu.h HINTS ${STARPU_DIR}/include/starpu/1.4 ${PC_STARPU_INCLUDE_DIRS})
find_library(STARPU_LIBRARY NAMES starpu-1.4 HINTS ${STARPU_DIR}/lib ${PC_STARPU_LIBRARY_DIRS})
include(FindPackageHandleStandardArgs)
find_package_handle_standard_args(StarPU REQUIRED_VARS STARPU_LIBRARY STARPU_INCLUDE_DIR)
if(StarPU_FOUND AND NOT TARGET StarPU::starpu)
  add_library(StarPU::starpu SHARED IMPORTED GLOBAL)
  set_target_properties(StarPU::starpu PROPERTIES IMPORTED_LOCATION "${STARPU_LIBRARY}"
                                                  INTERFACE_INCLUDE_DIRECTORIES "${STARPU_INCLUDE_DIR}")
endif()
