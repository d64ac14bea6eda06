/*
 * CPU restatement of the reference's batch assembly (host gather) and result
 * slicing (host scatter).  TEST INFRASTRUCTURE ONLY -- see oracle/__init__.py.
 * Built by oracle/Makefile into oracle/_build/liboracle_gs.so and called from
 * tests through ctypes.
 *
 * Byte-exact work: every function moves bytes and counts elements, nothing else.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#define ORACLE_OK 0
#define ORACLE_ERR_CAPACITY 1  /* InputPoolCapacityException */
#define ORACLE_ERR_ARG 2

/*
 * Host gather of ONE model input across the sub-requests of a batch.
 * Follows SlotManager::copy_job_inputs_to_slot
 * (/root/reference/src/starpu_task_worker/slot_manager_component.cpp:649-727):
 * the master job's tensor first, then each pending sub-job's tensor in order,
 * copied back to back at `offset += nbytes`; a copy that would pass the slot
 * capacity raises InputPoolCapacityException (:684-689); afterwards the StarPU
 * vector handle is resized to nx = sum(numel) (:722-724 ->
 * starpu_vector_resize_utils.hpp:19-64).
 * Known answer (reference test
 * tests/unit/starpu/unit_starpu_task_runner_submission_and_slots.cpp:607-666):
 * jobs {1,2} + {3,4} -> slot bytes 1,2,3,4 and nx == 4.
 */
int oracle_gather_concat(const void *const *srcs, const size_t *nbytes,
                         const size_t *numel, size_t n_jobs, void *slot,
                         size_t slot_capacity, size_t *nx_out,
                         size_t *bytes_out)
{
    size_t offset = 0, total_numel = 0;
    if ((!srcs || !nbytes || !numel || !slot) && n_jobs) return ORACLE_ERR_ARG;
    for (size_t j = 0; j < n_jobs; ++j) {
        if (!srcs[j]) continue; /* null pending job is skipped (:708-710) */
        if (slot_capacity < offset + nbytes[j]) return ORACLE_ERR_CAPACITY;
        memcpy((char *)slot + offset, srcs[j], nbytes[j]);
        offset += nbytes[j];
        total_numel += numel[j];
    }
    if (nx_out) *nx_out = total_numel;
    if (bytes_out) *bytes_out = offset;
    return ORACLE_OK;
}

/*
 * Ragged gather + pad + mask (the "variable-length admission" extension,
 * SURVEY.md section 8(f)-1; north_star stage 1).  Request r carries len[r]
 * token ids (and optionally len[r] mask words); row r of the batch is
 *   ids [r, 0:len)  = src_ids[r][0:len)      ids [r, len:S)  = pad_id
 *   mask[r, 0:len)  = src_mask[r] ? src_mask[r][0:len) : 1   mask[r, len:S) = 0
 * With every len[r] == S and one row per request this is exactly
 * oracle_gather_concat applied to the ids and to the mask input.
 */
int oracle_gather_pad_mask(const int64_t *const *src_ids,
                           const int64_t *const *src_mask, const int32_t *len,
                           size_t n_req, size_t seq, int64_t pad_id,
                           int64_t *ids_out, int64_t *mask_out)
{
    for (size_t r = 0; r < n_req; ++r) {
        if (len[r] < 0 || (size_t)len[r] > seq) return ORACLE_ERR_CAPACITY;
        for (size_t t = 0; t < seq; ++t) {
            const int in = t < (size_t)len[r];
            ids_out[r * seq + t] = in ? src_ids[r][t] : pad_id;
            mask_out[r * seq + t] =
                in ? ((src_mask && src_mask[r]) ? src_mask[r][t] : 1) : 0;
        }
    }
    return ORACLE_OK;
}

/*
 * Host scatter: slice [offset, offset+batch_size) along dim 0 of an aggregated
 * contiguous output and copy it to the sub-request's buffer.
 * Follows slice_outputs_for_sub_job / slice_contiguous_output_for_sub_job
 * (/root/reference/src/starpu_task_worker/batching_helpers.hpp:74-127, 3-46):
 * slice_size = max(1, batch_size); end clamped to the available rows; length<=0
 * yields an empty output; sample_bytes = total_bytes / rows.  The reference
 * returns a zero-copy view; the bytes that view exposes are what is written to
 * `dst` here.  Returns the processed length (rows) through *rows_out.
 * Known answer (reference test
 * tests/unit/starpu/unit_starpu_task_runner_batching_and_internal.cpp:1514-1606):
 * batch sizes 1+2 over 3 rows -> rows [0,1) and [1,3).
 */
int oracle_scatter_slice(const void *aggregated, size_t rows,
                         size_t total_bytes, int64_t offset, int64_t batch_size,
                         void *dst, int64_t *rows_out)
{
    const int64_t slice_size = batch_size > 1 ? batch_size : 1;
    int64_t end, length;
    size_t sample_bytes;
    if (rows_out) *rows_out = 0;
    if (!aggregated || rows == 0 || offset < 0) return ORACLE_OK;
    if (total_bytes % rows != 0) return ORACLE_ERR_ARG;
    end = offset + slice_size;
    if (end > (int64_t)rows) end = (int64_t)rows;
    length = end - offset;
    if (length <= 0) return ORACLE_OK;
    sample_bytes = total_bytes / rows;
    memcpy(dst, (const char *)aggregated + (size_t)offset * sample_bytes,
           (size_t)length * sample_bytes);
    if (rows_out) *rows_out = length;
    return ORACLE_OK;
}

/*
 * Ragged scatter (pairs with oracle_gather_pad_mask): request r receives the
 * first len[r] token rows of batch row r, each `row_bytes` wide.
 */
int oracle_scatter_ragged(const void *aggregated, size_t n_req, size_t seq,
                          size_t row_bytes, const int32_t *len,
                          void *const *dsts)
{
    for (size_t r = 0; r < n_req; ++r) {
        if (len[r] < 0 || (size_t)len[r] > seq) return ORACLE_ERR_CAPACITY;
        memcpy(dsts[r], (const char *)aggregated + r * seq * row_bytes,
               (size_t)len[r] * row_bytes);
    }
    return ORACLE_OK;
}

/*
 * buffer_byte_size (/root/reference/src/core/starpu_setup.cpp:515-542):
 * VARIABLE -> elemsize; VECTOR -> nx*elemsize with overflow check.
 * interface ids: STARPU_VECTOR_INTERFACE_ID = 2, STARPU_VARIABLE_INTERFACE_ID = 5
 * (StarPU 1.4 enum starpu_data_interface_id: MATRIX 0, BLOCK 1, VECTOR 2, CSR 3,
 * BCSR 4, VARIABLE 5, ...).
 */
int oracle_buffer_byte_size(int interface_id, uint64_t nx, uint64_t elemsize,
                            uint64_t *out)
{
    if (interface_id == 5) { *out = elemsize; return ORACLE_OK; }
    if (interface_id == 2) {
        if (elemsize != 0 && nx > UINT64_MAX / elemsize) return ORACLE_ERR_CAPACITY;
        *out = nx * elemsize;
        return ORACLE_OK;
    }
    return ORACLE_ERR_ARG;
}
