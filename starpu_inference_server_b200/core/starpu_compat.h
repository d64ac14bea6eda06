/*
 * starpu_compat.h -- the slice of the StarPU 1.4 C API the inference hot path touches.
 *
 * The reference links the real runtime (cmake/FindStarPU.cmake:10-28).  StarPU is not
 * installed in this image (no starpu.h / libstarpu-1.4), so when <starpu.h>
 * is absent this header declares the same types, constants and functions with the same
 * names and field names, and core/mini_starpu.cpp implements them with a small in-process
 * runtime (one thread + one stream per CUDA worker, a central eager priority queue, MSI-
 * style host<->device coherence per vector handle).  The codelet source
 * (core/inference_codelet.cpp) is written against these names only, so it compiles
 * unchanged against real StarPU: build with -DSPU_USE_REAL_STARPU and the include path.
 *
 * API list = SURVEY.md 8(b) last row (grep of /root/reference/src).
 */
#ifndef SPU_STARPU_COMPAT_H
#define SPU_STARPU_COMPAT_H

#if defined(SPU_USE_REAL_STARPU)
#include <starpu.h>
#include <starpu_cuda.h>
#else

#include <limits.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define STARPU_MAIN_RAM 0
#define STARPU_MAXIMPLEMENTATIONS 4
#define STARPU_NMAXBUFS 8
#define STARPU_NMAXWORKERS 128
#define STARPU_MAXNODES 16
#define STARPU_VARIABLE_NBUFFERS (-1)
#define STARPU_CUDA_ASYNC (1 << 0)
#define STARPU_DEFAULT_PRIO 0
#define STARPU_CPU (1ULL << 1)
#define STARPU_CUDA (1ULL << 3)

enum starpu_data_access_mode { STARPU_NONE = 0, STARPU_R = (1 << 0), STARPU_W = (1 << 1), STARPU_RW = 3 };
enum starpu_codelet_type { STARPU_SEQ = 0, STARPU_SPMD, STARPU_FORKJOIN };
enum starpu_worker_archtype { STARPU_CPU_WORKER = 0, STARPU_CUDA_WORKER = 1, STARPU_ANY_WORKER = 255 };
/* StarPU 1.4 enum starpu_data_interface_id */
enum starpu_data_interface_id {
  STARPU_UNKNOWN_INTERFACE_ID = -1,
  STARPU_MATRIX_INTERFACE_ID = 0,
  STARPU_BLOCK_INTERFACE_ID = 1,
  STARPU_VECTOR_INTERFACE_ID = 2,
  STARPU_CSR_INTERFACE_ID = 3,
  STARPU_BCSR_INTERFACE_ID = 4,
  STARPU_VARIABLE_INTERFACE_ID = 5
};

struct starpu_vector_interface {
  enum starpu_data_interface_id id;
  uintptr_t ptr;        /* local pointer of the vector on this memory node */
  uintptr_t dev_handle;
  size_t offset;
  size_t allocsize;     /* bytes really allocated (slot capacity) */
  uint32_t nx;          /* number of elements */
  size_t elemsize;
  uint32_t slice_base;
};

struct starpu_variable_interface {
  enum starpu_data_interface_id id;
  uintptr_t ptr;
  uintptr_t dev_handle;
  size_t offset;
  size_t elemsize;
};

typedef struct _starpu_data_state *starpu_data_handle_t;
typedef void (*starpu_cpu_func_t)(void **, void *);
typedef void (*starpu_cuda_func_t)(void **, void *);

struct starpu_codelet {
  uint32_t where;
  enum starpu_codelet_type type;
  int max_parallelism;
  starpu_cpu_func_t cpu_funcs[STARPU_MAXIMPLEMENTATIONS];
  starpu_cuda_func_t cuda_funcs[STARPU_MAXIMPLEMENTATIONS];
  char cuda_flags[STARPU_MAXIMPLEMENTATIONS];
  int nbuffers;
  enum starpu_data_access_mode modes[STARPU_NMAXBUFS];
  enum starpu_data_access_mode *dyn_modes;
  const char *name;
};

struct starpu_task {
  const char *name;
  struct starpu_codelet *cl;
  int nbuffers;
  starpu_data_handle_t handles[STARPU_NMAXBUFS];
  starpu_data_handle_t *dyn_handles;
  enum starpu_data_access_mode modes[STARPU_NMAXBUFS];
  enum starpu_data_access_mode *dyn_modes;
  void *cl_arg;
  size_t cl_arg_size;
  void (*callback_func)(void *);
  void *callback_arg;
  int priority;
  unsigned synchronous : 1;
  unsigned destroy : 1;
  unsigned execute_on_a_specific_worker : 1;
  unsigned workerid;
  void *starpu_private;
};

struct starpu_conf {
  const char *sched_policy_name;
  int ncpus;
  int ncuda;
  unsigned use_explicit_workers_cuda_gpuid;
  unsigned workers_cuda_gpuid[STARPU_NMAXWORKERS];
  /* mini-runtime only: real StarPU reads STARPU_NWORKER_PER_CUDA / STARPU_CUDA_PIPELINE
   * from the environment (models/bert.yml:6-8); 0 = take them from the environment too. */
  int nworker_per_cuda;
  int cuda_pipeline;
};

int starpu_conf_init(struct starpu_conf *conf);
int starpu_init(struct starpu_conf *conf);
void starpu_shutdown(void);
int starpu_is_initialized(void);

void starpu_codelet_init(struct starpu_codelet *cl);
struct starpu_task *starpu_task_create(void);
int starpu_task_submit(struct starpu_task *task);
void starpu_task_destroy(struct starpu_task *task);
int starpu_task_wait_for_all(void);

void starpu_vector_data_register(starpu_data_handle_t *handle, int home_node, uintptr_t ptr,
                                 uint32_t nx, size_t elemsize);
void starpu_variable_data_register(starpu_data_handle_t *handle, int home_node, uintptr_t ptr,
                                   size_t size);
void starpu_data_unregister(starpu_data_handle_t handle);
void starpu_data_unregister_submit(starpu_data_handle_t handle);
int starpu_data_acquire(starpu_data_handle_t handle, enum starpu_data_access_mode mode);
int starpu_data_acquire_cb(starpu_data_handle_t handle, enum starpu_data_access_mode mode,
                           void (*callback)(void *), void *arg);
void starpu_data_release(starpu_data_handle_t handle);
/* write-through: every write to the handle is propagated to the memory nodes whose bit is set */
void starpu_data_set_wt_mask(starpu_data_handle_t handle, uint32_t wt_mask);
enum starpu_data_interface_id starpu_data_get_interface_id(starpu_data_handle_t handle);
void *starpu_data_get_interface_on_node(starpu_data_handle_t handle, unsigned node);
unsigned starpu_memory_nodes_get_count(void);
int starpu_memory_pin(void *addr, size_t size);
int starpu_memory_unpin(void *addr, size_t size);

int starpu_worker_get_id(void);
int starpu_worker_get_devid(int id);
unsigned starpu_worker_get_count(void);
enum starpu_worker_archtype starpu_worker_get_type(int id);
unsigned starpu_worker_get_ids_by_type(enum starpu_worker_archtype type, int *workerids,
                                       unsigned maxsize);
unsigned starpu_worker_get_memory_node(unsigned workerid);
int starpu_sched_get_min_priority(void);
int starpu_sched_get_max_priority(void);
/* cudaStream_t; declared void* here so that this header needs no CUDA include */
void *starpu_cuda_get_local_stream_ptr(void);
#define starpu_cuda_get_local_stream() ((cudaStream_t)starpu_cuda_get_local_stream_ptr())

#ifdef __cplusplus
}
#endif
#endif /* SPU_USE_REAL_STARPU */
#endif /* SPU_STARPU_COMPAT_H */
