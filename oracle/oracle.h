/* TEST INFRASTRUCTURE ONLY — C entry points of the CPU oracle (liboracle.so).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this library.  Struct layouts are shared with include/trajopt_cuda.h so the checker and the product
 * can be compared field by field; the implementation is independent (oracle/oracle.cpp). */
#ifndef ORACLE_H
#define ORACLE_H
#include "../include/trajopt_cuda.h"
#ifdef __cplusplus
extern "C" {
#endif
const char* orc_last_error(void);
int orc_scene_create(const tc_scene_desc* desc, void** out);
void orc_scene_destroy(void* scene);
int orc_scene_num_pairs(const void* scene);
int orc_scene_pairs(const void* scene, int32_t* l0, int32_t* l1, double* margin, double* coeff);
int orc_fk(const void* scene, const double* q, double* link_poses);
void orc_sincos(double x, double* sn, double* cs);
int orc_contact_test(void* scene, const double* pose0, const double* pose1, int32_t n_states, int32_t type, void** out);
int orc_jacobian(const void* scene, const double* q, int link, const double* ref_offset, double* J);
int orc_shape_distance(const void* scene, int shape_a, const double* pose_a0, const double* pose_a1, int shape_b,
                       const double* pose_b0, const double* pose_b1, double threshold, int* contact, double* distance,
                       double* normal, double* pa, double* pb, int* used_epa, int* iters);
int orc_shape_support(const void* scene, int shape, const double* pose0, const double* pose1, const double* dir,
                      double* value);
int orc_eval(void* scene, const tc_eval_desc* desc, const double* q, int32_t B, int32_t T, void** out, int n_threads);
void orc_result_destroy(void* res);
int orc_result_counts(const void* res, tc_counts* c);
int orc_result_contacts(const void* res, tc_contact* dst, size_t cap);
int orc_result_contact_transforms(const void* res, double* dst, size_t cap);
int orc_result_term_ptr(const void* res, int64_t* term_ptr);
int orc_result_gradients(const void* res, tc_gradient* rec, double* grad, size_t cap);
int orc_result_rows(const void* res, int64_t* row_ptr, int32_t* col_idx, double* val, double* row_value,
                    double* row_coeff, double* row_margin);
int orc_result_sets(const void* res, int32_t max_num_cnt, int64_t* set_ptr, tc_set* sets, size_t cap_sets,
                    double* values, double* coeffs, double* jac);
int orc_result_cost(const void* res, double* cost_per_term, double* cost_per_traj);
int orc_num_threads(void);
#ifdef __cplusplus
}
#endif
#endif
