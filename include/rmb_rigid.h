/*
 * rmb_rigid.h -- C ABI of the rigid-body operators around the blob mobility
 * product (librmb_mobility.so).  They replace interpreted Python loops over
 * bodies that run inside every GMRES matvec of the reference:
 *
 *   rmb_blobs_from_bodies     body/body.py:64-80          Body.get_r_vectors
 *   rmb_K_apply               multi_bodies.py:328-347     K_matrix_vector_prod
 *   rmb_KT_apply              multi_bodies.py:350-371     K_matrix_T_vector_prod
 *   rmb_body_mobility_blocks  multi_bodies.py:772-774 + mobility.py:1018
 *                             (Body.calc_mobility_blobs -> single_wall_fluid_mobility)
 *
 * Same conventions as rmb_mobility.h: device pointers, float64, enqueue on
 * `stream`, never synchronise, 0 on success.  Bodies own contiguous blob
 * ranges: body b owns blobs [body_offset[b], body_offset[b+1]).
 */
#ifndef RMB_RIGID_H
#define RMB_RIGID_H

#ifdef __cplusplus
extern "C" {
#endif

const char *rmb_rigid_last_error(void);

/* r_i = R(quaternion_b) ref_i + location_b.  quaternion = (s, p1, p2, p3) as quaternion.py:6-10.
 *   blob_body   device int32 [n_blobs]    index of the body that owns blob i
 *   ref_config  device [n_blobs][3]       Body.reference_configuration, concatenated
 *   location    device [n_bodies][3], quaternion device [n_bodies][4], r_out device [n_blobs][3] */
int rmb_blobs_from_bodies(long n_blobs, const int *blob_body, const double *ref_config,
                          const double *location, const double *quaternion, double *r_out, void *stream);

/* out_i = beta*out_i + alpha*(U_b + omega_b x (r_i - location_b)),  U = [n_bodies][6] = (U, omega) */
int rmb_K_apply(long n_blobs, const int *blob_body, const double *r, const double *location,
                const double *U, double alpha, double beta, double *out, void *stream);

/* out_b = alpha * [ sum_i lambda_i ; sum_i (r_i - location_b) x lambda_i ],  out = [n_bodies][6]
 *   body_offset device int64 [n_bodies + 1] */
int rmb_KT_apply(long n_bodies, const long *body_offset, const double *r, const double *location,
                 const double *lambda, double alpha, double *out, void *stream);

/* Dense blob mobility (UF block, with the wall clamp + damping when wall = 1) of every body:
 * M = [n_bodies][3 nb_max][3 nb_max] row-major, identity-padded for bodies with fewer than nb_max blobs. */
int rmb_body_mobility_blocks(long n_bodies, const long *body_offset, int nb_max, const double *r,
                             double eta, double a, int wall, double *M, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* RMB_RIGID_H */
