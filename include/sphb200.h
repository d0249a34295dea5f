/* sphb200.h — C ABI of the B200-native SPHinXsys particle-interaction path.
 *
 * Plain C: opaque handles, plain pointers and sizes, no C++/torch types.  One context per GPU
 * (per rank).  Every entry point cites the reference interface it replaces; paths are relative to
 * the reference tree's src/shared/ unless noted.  All calls are asynchronous on the context's
 * stream except those that return a value to the host (…_dt with a non-NULL out pointer,
 * var_to_host, reductions, debug dumps), which synchronise that stream.
 *
 * Error convention: the reference prints and exit(1)s (e.g. particle_iterators.h:51-53); this ABI
 * returns a status and keeps a per-context message (sphb200_last_error).  It never exits.
 *
 * Ordering contract: sphb200_cell_list_update physically reorders the body's particles by linear
 * cell index (the cell list IS the counting sort), permuting every registered variable and keeping
 * OriginalID / SortedID consistent exactly as ParticleSorting does
 * (particle_dynamics/configuration_dynamics/particle_sorting.cpp:16-23,55-58).
 */
#ifndef SPHB200_H
#define SPHB200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct sphb200_ctx sphb200_ctx;
typedef struct sphb200_body sphb200_body;
typedef struct sphb200_relation sphb200_relation;
typedef int sphb200_status;

enum
{
    SPHB200_OK = 0,
    SPHB200_ERR_INVALID = 1,   /* bad argument / call order */
    SPHB200_ERR_CUDA = 2,      /* CUDA runtime error (message holds cudaGetErrorString) */
    SPHB200_ERR_NOT_FOUND = 3, /* variable not registered */
    SPHB200_ERR_NO_DEVICE = 4  /* no CUDA device: there is NO CPU fallback */
};

/* ---- context ------------------------------------------------------------------------------- */
/* Replaces SPHSystem (sphinxsys_system/sph_system.cpp:36-51) for the path: system bounds are the
 * already expanded bounds (user box ± 4 dp).  ndim = 2|3 (SPH_NDIM), real_bytes = 8|4
 * (SPHINXSYS_USE_FLOAT, common/base_data_type.h:50-60). */
sphb200_status sphb200_ctx_create(int device, int ndim, int real_bytes, const double *sys_lo, const double *sys_hi,
                                  sphb200_ctx **out);
void sphb200_ctx_destroy(sphb200_ctx *ctx);
sphb200_status sphb200_ctx_synchronize(sphb200_ctx *ctx);
const char *sphb200_last_error(sphb200_ctx *ctx);
/* the cudaStream_t all work of this context is enqueued on (for CUDA-event timing by the caller) */
void *sphb200_ctx_stream(sphb200_ctx *ctx);
/* number of kernels this context has launched so far (bench.py's gpu_launches) */
int64_t sphb200_ctx_launch_count(sphb200_ctx *ctx);
/* which pair-sum kernels the acoustic sub-step uses: -1 automatic (default: group-tiled kernels on union
 * neighbour lists, list-per-particle kernels for periodic bodies), 0 always
 * list-per-particle, 1 group-tiled wherever the lists allow it.  Both give the reference's sums; the knob exists
 * for A/B timing and for the parity tests of both paths (env SPHB200_TILE sets the process-wide default). */
sphb200_status sphb200_ctx_set_pair_kernel(sphb200_ctx *ctx, int mode);

/* ---- bodies and named variables ------------------------------------------------------------ */
/* SPHBody/RealBody + SPHAdaptation (bodies/base_body.h:62-201, adaptations/adaptation.cpp:12-19):
 * h = h_ratio*dp, Wendland C2 (kernels/kernel_wendland_c2.cpp:8-35), cell-linked-list mesh with
 * spacing = cut-off radius and 2 buffer cells (meshes/cell_linked_list.cpp:10-20, base_mesh.cpp:6-16),
 * sigma0 = lattice number density (adaptation.cpp:26-60). */
sphb200_status sphb200_body_create(sphb200_ctx *ctx, const char *name, int64_t n, double dp, double h_ratio,
                                   sphb200_body **out);
/* The spacing of the body's cell-linked-list mesh.  The reference builds every body's cell list with the body's OWN cut-off
 * radius (meshes/cell_linked_list.cpp:84-101 through its adaptation) and lets a contact search from a body with a larger
 * kernel walk a deeper block of cells (SearchDepthContact, body_relations/base_body_relation.cpp:10-16); the accepted
 * neighbour set does not depend on the target's cell size.  A contact relation whose targets have cells at least as large
 * as the SOURCE's cut-off takes the single-pass masked search here, so a caller may widen a target's cells (a SolidBody
 * wall with h = 1.15 dp searched by a fluid with h = 1.3 dp: spacing = 2.6 dp).  Invalidates the cell list and every
 * relation of the body: call cell_list_update / relation_update afterwards. */
sphb200_status sphb200_body_set_cell_spacing(sphb200_body *b, double spacing);
/* WeaklyCompressibleFluid(rho0,c0) + Viscosity(mu) (materials/weakly_compressible_fluid.cpp:8-30) */
sphb200_status sphb200_body_set_fluid(sphb200_body *b, double rho0, double c0, double mu);
/* NeoHookeanSolid(rho0,E,nu) (materials/elastic_solid.cpp:54-80) */
sphb200_status sphb200_body_set_elastic(sphb200_body *b, double rho0, double youngs, double poisson);
/* PeriodicAlongAxis (particle_dynamics/general_dynamics/domian_bouding/domain_bounding.h:49-70) */
sphb200_status sphb200_body_set_periodic(sphb200_body *b, int axis, double lo, double hi);
/* out[0..11] = h, rc, W0, sigma0, mesh lo[3], mesh spacing, cells[3], n */
sphb200_status sphb200_body_info(sphb200_body *b, double *out12);

/* BaseParticles::registerStateVariable / addEvolvingVariable (particles/base_particles.h:119-162).
 * width = Reals per particle (1 scalar, ndim Vecd, ndim*ndim Matd), array-of-Vecd layout
 * (common/sphinxsys_variable.h:204-236).  host_ptr is BORROWED for the body's lifetime (may be NULL
 * for a device-only variable, cf. DeviceOnlyDiscreteVariable, sphinxsys_variable.h:365-377).
 * The device mirror is zero-initialised. Registering an existing name rebinds host_ptr. */
sphb200_status sphb200_var_register(sphb200_body *b, const char *name, int width, int evolving, void *host_ptr);
/* DiscreteVariable::synchronizeToDevice / synchronizeWithDevice (sphinxsys_variable.h:179-190;
 * SYCL twin src_sycl/shared/common/sphinxsys_variable_sycl.hpp:35-87) */
sphb200_status sphb200_var_to_device(sphb200_body *b, const char *name);
sphb200_status sphb200_var_to_host(sphb200_body *b, const char *name);
/* explicit-buffer variants (host buffer of n*width Reals) */
sphb200_status sphb200_var_upload(sphb200_body *b, const char *name, const void *src);
sphb200_status sphb200_var_download(sphb200_body *b, const char *name, void *dst);
/* Double-buffered upload.  var_stage starts an asynchronous copy of a host buffer (n*width Reals; pinned memory
 * for a real overlap) into a staging mirror on the context's copy stream - the dynamics queued on the compute
 * stream keep running on the current mirrors.  The host buffer stays RESERVED until sphb200_ctx_stage_wait (or
 * sphb200_ctx_synchronize) returns: body_commit_staged only orders the compute stream after the copies, it does not
 * block the host.  body_commit_staged swaps every staged mirror in (like var_upload, the variables count as
 * rewritten: call cell_list_update afterwards).  If the body was re-sorted (cell_list_update) between var_stage and
 * the commit, the staged rows are in another particle order than the resident ones: the commit then requires EVERY
 * evolving variable to be staged (status INVALID otherwise), the staged row order becomes the body's order,
 * OriginalID/SortedID restart from the identity (sphb200_ids_from_host restores saved ones) and the cell list and
 * relations are invalidated - the semantics of a reload (io_system/io_base.cpp:67-97, Dambreak.cpp:132-138).
 * The SYCL twin is the queue-ordered copy of src_sycl/shared/common/sphinxsys_variable_sycl.hpp:35-87, which the
 * host thread does not wait for either. */
sphb200_status sphb200_var_stage(sphb200_body *b, const char *name, const void *src);
sphb200_status sphb200_body_commit_staged(sphb200_body *b);
sphb200_status sphb200_ctx_stage_wait(sphb200_ctx *ctx);
/* Asynchronous, double-buffered device -> host read for the output / restart path
 * (BaseVariable::prepareForOutput(ParallelDevicePolicy), common/sphinxsys_variable.h:350-356;
 * BodyStatesRecording / RestartIO::writeToFile, io_system/io_base.cpp:67-97): a device-side snapshot is taken in
 * stream order, the copy to dst (n*width Reals, pinned for a real overlap) runs on a separate stream while the
 * dynamics continue; rows are in the body's order at the time of the call.  ctx_fetch_wait blocks the host until
 * every fetch issued so far has landed. */
sphb200_status sphb200_var_fetch_async(sphb200_body *b, const char *name, void *dst);
sphb200_status sphb200_ctx_fetch_wait(sphb200_ctx *ctx);
/* DelegatedData(par_device): raw device pointer, valid until the next cell_list_update of b */
sphb200_status sphb200_var_device_ptr(sphb200_body *b, const char *name, void **out);
/* Reals per particle of a registered variable (DiscreteVariable<T>: sizeof(T)/sizeof(Real),
 * sphinxsys_variable.h:196-236); dynamics register their own outputs (e.g. "ViscousForce") on first use */
sphb200_status sphb200_var_width(sphb200_body *b, const char *name, int *width);
/* OriginalID / SortedID (particles/base_particles.cpp:33-38), widened to int64 */
sphb200_status sphb200_ids_to_host(sphb200_body *b, int64_t *original_id, int64_t *sorted_id);
/* restores OriginalID (a permutation of 0..n-1, current row order) after a reload and rebuilds SortedID
 * (particle_sorting.cpp:55-58; the reference's restart file carries OriginalID, base_particles.cpp:33-37) */
sphb200_status sphb200_ids_from_host(sphb200_body *b, const int64_t *original_id);

/* ---- spatial search ------------------------------------------------------------------------ */
/* RealBody::updateCellLinkedList (bodies/base_body.cpp:113-116 → meshes/cell_linked_list.cpp:84-101;
 * CK twin shared_ck/.../update_cell_linked_list.hpp:75-106) fused with ParticleSorting::exec
 * (configuration_dynamics/particle_sorting.hpp:15-21): histogram → scan → scatter, then every
 * variable is gathered into cell order.  Takes the position snapshot the pair geometry is frozen to. */
sphb200_status sphb200_cell_list_update(sphb200_body *b);
/* parity/debug: per particle (current order) linear cell index (base_mesh.hxx:68-78) and the
 * reference's 10-bit Morton key (base_mesh.hxx:80-99) */
sphb200_status sphb200_cell_keys(sphb200_body *b, int64_t *linear, int64_t *morton);
/* parity/debug: CellOffset array (cells+1 entries) of the CSR cell list */
sphb200_status sphb200_cell_offsets(sphb200_body *b, int64_t *out, int64_t *n_cells);
/* PeriodicConditionUsingCellLinkedList::bounding_ (domain_bounding.h:93-137): wrap positions */
sphb200_status sphb200_periodic_bounding(sphb200_body *b, int axis);

/* InnerRelation / ContactRelation (body_relations/inner_body_relation.cpp:14-20,
 * contact_body_relation.cpp:22-32) */
sphb200_status sphb200_relation_inner_create(sphb200_body *b, sphb200_relation **out);
sphb200_status sphb200_relation_contact_create(sphb200_body *b, sphb200_body *const *targets, int n_targets,
                                               sphb200_relation **out);
/* The solid-contact neighbour builders on the same cell lists.  part_original_ids (n_part > 0) restricts
 * the SOURCE particles to a body part by particle (the reference's BodySurfaceLayer; its geometry is the caller's),
 * the other particles get empty neighbourhoods.
 *   self contact   : SelfSurfaceContactRelation + NeighborBuilderSelfContact (inner_body_relation.cpp:50-75,
 *                    neighborhood.cpp:123-142): j is a neighbour iff |x_i - x_j| < rc and |X_i - X_j| > rc, X = the
 *                    body's "InitialPosition" variable
 *   surface contact: SurfaceContactRelation + NeighborBuilderSurfaceContact (contact_body_relation.cpp:192-240,
 *                    neighborhood.cpp:162-168): the contact builder with the Wendland kernel of the AVERAGED smoothing
 *                    length, cut-off 2 * 0.5 * (h_source + h_target) */
sphb200_status sphb200_relation_self_contact_create(sphb200_body *b, const int64_t *part_original_ids, int64_t n_part, sphb200_relation **out);
sphb200_status sphb200_relation_surface_contact_create(sphb200_body *b, sphb200_body *const *targets, int n_targets,
                                                       const int64_t *part_original_ids, int64_t n_part, sphb200_relation **out);
/* SPHRelation::updateConfiguration (base_body_relation.h:100; searchNeighborsByMesh
 * meshes/cell_linked_list.hpp:20-46 with NeighborBuilderInner/Contact neighborhood.cpp:85-98,147-160).
 * Builds a compact u32 neighbour-index list (count → scan → fill, cf. CK update_body_relation.hpp:117-164). */
sphb200_status sphb200_relation_update(sphb200_relation *r);
/* parity/debug: counts[n] and flat neighbour indices (list order) of target k; returns the total */
sphb200_status sphb200_relation_neighbors(sphb200_relation *r, int k, int32_t *counts, int64_t *j_out, int64_t *total);
/* measurement: out8 = {accepted pairs, 1 if the single-pass masked search held (0: a candidate run overflowed its
 * mask words or the target is periodic -> generic two-pass search), 1 if the u32 index list is materialised,
 * union-list entries, union groups, particles per group (0 until a group-tiled kernel built it), n_source, build id} */
sphb200_status sphb200_relation_stats(sphb200_relation *r, int k, double *out8);

/* ---- weakly compressible fluid dynamics ---------------------------------------------------- */
/* DensitySummation<Inner<>|Inner<FreeSurface>> (+Contact<>) interaction+update
 * (fluid_dynamics/density_summation.cpp:8-22,58-78; density_summation.hpp:11-32) */
sphb200_status sphb200_density_summation(sphb200_relation *inner, sphb200_relation *contact_or_null, int free_surface);
/* Integration1stHalf<Inner<>,Riemann|NoRiemann,NoKernelCorrection> (+Contact<Wall>) as Dynamics1Level:
 * initialization + interaction + update (fluid_dynamics/fluid_integration.hpp:50-113).  Also
 * accumulates max(c0+|v|) on the device so the following sphb200_acoustic_dt costs no extra pass. */
sphb200_status sphb200_integration_1st_half(sphb200_relation *inner, sphb200_relation *wall_or_null, int riemann, double dt);
/* Integration2ndHalf<Inner<>,Riemann|NoRiemann> (+Contact<Wall>) (fluid_integration.hpp:167-231) */
sphb200_status sphb200_integration_2nd_half(sphb200_relation *inner, sphb200_relation *wall_or_null, int riemann, double dt);
/* ReduceDynamics<AcousticTimeStep> (fluid_dynamics/fluid_time_step.cpp:12-31): cfl*h/(max(c0+|v|)+tiny) */
sphb200_status sphb200_acoustic_dt(sphb200_body *b, double cfl, double *dt_out);
/* ReduceDynamics<AdvectionTimeStep|AdvectionViscousTimeStep> (fluid_time_step.cpp:44-79).  dt_out == NULL: the
 * result stays on the device (the next sphb200_acoustic_relaxation with Dt < 0 uses it), no host synchronisation. */
sphb200_status sphb200_advection_dt(sphb200_body *b, double U_ref, double cfl, int viscous, double *dt_out);
/* The whole acoustic sub-step loop of one advection step,
 *     while (relaxation_time < Dt) { Integration1stHalf; Integration2ndHalf; dt = AcousticTimeStep; ... }
 * with dt, relaxation_time and the exit test resident on the device: every sub-step kernel reads dt from a control
 * block and returns at once after the loop has ended, so the host queues sub-steps without waiting for the
 * ReduceDynamics result (the reference's particle_reduce returns to the host every time,
 * particle_iterators.h / src_sycl .../particle_iterators_sycl.h:80-105).  order selects the reference main's statement order:
 *   0  dt = reduce; step(dt); relaxation_time += dt                  tests/2d_examples/test_2d_dambreak/Dambreak.cpp:183-192
 *   1  step(dt of the previous sub-step); dt = reduce; relaxation_time += dt      tests/3d_examples/test_3d_dambreak/dambreak.cpp:168-177
 *   2  dt = min(reduce, Dt); relaxation_time += dt; step(dt)         tests/2d_examples/test_2d_taylor_green/taylor_green.cpp:167-176
 * Dt < 0: use the device-resident Dt of the last sphb200_advection_dt(..., NULL).  dt_carry >= 0 sets the dt the first
 * sub-step integrates with (order 1), < 0 keeps the device's.  out5 != NULL: sub-steps are queued max_substeps at first,
 * then 4 at a time until the device reports the end of the loop (one host synchronisation per batch) and
 * out5 = {sub-steps done, last dt, accumulated physical time, relaxation_time, 1}.  out5 == NULL: exactly max_substeps
 * launches are queued and the call returns without synchronising; sphb200_relaxation_state reads the block later
 * (out5[4] = 0 there: the loop needs more sub-steps - call again with resume = 1). */
sphb200_status sphb200_acoustic_relaxation(sphb200_relation *inner, sphb200_relation *wall_or_null, int riemann_1st, int riemann_2nd,
                                           double Dt, double cfl, int order, double dt_carry, int max_substeps, int resume, double *out5);
sphb200_status sphb200_relaxation_state(sphb200_ctx *ctx, double *out5);
/* The same loop in pieces, for a driver that interleaves its own work with the stages (the slab run: halo records
 * between the stages, an all-reduce of the max slot over the ranks before the bookkeeping): relaxation_begin sets Dt
 * and the carried dt; the stage entry points (integration_1st_half[_init|_interact], integration_2nd_half) called
 * with dt < 0 read dt and the exit flag from the control block; substep_ctl(phase 0 | 1) is the before / after
 * bookkeeping of one sub-step (see sphb200_acoustic_relaxation for `order`); reduce_slot_ptr gives the device address
 * of reduction slot `slot` (0: bit pattern of max(c0 + |v|), a non-negative double - an integer MAX all-reduce of it
 * over the ranks is exact). */
sphb200_status sphb200_relaxation_begin(sphb200_ctx *ctx, double Dt, double dt_carry);
sphb200_status sphb200_substep_ctl(sphb200_body *b, double cfl, int order, int phase);
sphb200_status sphb200_reduce_slot_ptr(sphb200_ctx *ctx, int slot, void **out);
/* SimpleDynamics<GravityForce<Gravity>> (general_dynamics/force_prior.hpp:22-41) */
sphb200_status sphb200_gravity_force(sphb200_body *b, const double *g);
/* InteractionWithUpdate<ViscousForceInner> (fluid_dynamics/viscous_dynamics.hpp:35-53 + force_prior.hpp:22-26), or with
 * wall_or_null != NULL InteractionWithUpdate<ViscousForceWithWall> = inner + ViscousForce<Contact<Wall>>
 * (viscous_dynamics.hpp:92-115: 2 (v_i - v_wall) / (r + 0.01 h), weight 2 e.(B_i e) mu, added onto the inner force) */
sphb200_status sphb200_viscous_force(sphb200_relation *inner, sphb200_relation *wall_or_null);
/* InteractionWithUpdate<TransportVelocityCorrectionInner<…TruncatedLinear,AllParticles>>
 * (fluid_dynamics/transport_velocity_correction.hpp:38-67) */
sphb200_status sphb200_transport_velocity_correction(sphb200_relation *inner, double coefficient);
/* ReducedQuantityRecording observables (general_dynamics/general_reduce.cpp:18-79) */
sphb200_status sphb200_total_kinetic_energy(sphb200_body *b, double *out);
sphb200_status sphb200_total_mechanical_energy(sphb200_body *b, const double *g, double *out);
sphb200_status sphb200_maximum_speed(sphb200_body *b, double *out);
/* ReduceDynamics<PositionUpperBound | PositionLowerBound> (general_reduce.cpp:28-49): component-wise bound of the
 * particle positions, out3[ndim..2] = 0.  The dam-break front position is the upper bound's x component. */
sphb200_status sphb200_position_bound(sphb200_body *b, int upper, double *out3);
/* ObservingAQuantity (general_dynamics/general_interpolation.h:63-83): out = n_observer*width doubles */
sphb200_status sphb200_observe(sphb200_relation *contact, const char *name, int width, double *out);

/* ---- CK-flavour fluid dynamics of the reference's SYCL dam-break tests ------------------ */
/* All three evaluate W / dW with the CK smoothing kernel, the 20-interval cubic-Lagrange table of Wendland C2
 * (shared_ck/smoothing_kernel/kernel_tabulated_ck.h:37-71; inv_h = 1 / max(h_source, h_target), neighbor_method.hpp:73-120).
 * CompressionSummation<Inner, Contact> + DensityRegularization<FreeSurface | Internal>
 * (shared_ck/particle_dynamics/fluid_dynamics/density_regularization.hpp:42-111): CompressionSummation = W0 Vref_i + sum W_ij Vref_j,
 * Compression = max(sum, 1) | sum, Density = Compression * rho0; "VolumetricMeasureRef" is copied from VolumetricMeasure when first used */
sphb200_status sphb200_ck_density_regularization(sphb200_relation *inner, sphb200_relation *contact_or_null, int free_surface);
/* the same as its two reference objects: InteractionDynamicsCK<..., CompressionSummation<Inner<>, Contact<>>>::exec()
 * (density_regularization.hpp:42-83: the sums into "CompressionSummation") and StateDynamics<..., DensityRegularization<...,
 * FreeSurface | Internal>>::exec() (:103-111: Compression = max(sum, 1) | sum, Density = Compression rho0) */
sphb200_status sphb200_ck_compression_summation(sphb200_relation *inner, sphb200_relation *contact_or_null);
sphb200_status sphb200_ck_density_regularize(sphb200_body *body, int free_surface);
/* LinearCorrectionMatrix<Inner<WithUpdate>>(alpha) + <Contact> (general_dynamics/kernel_correction_ck.hpp:36-96): B = -sum r_ij (nablaW_ij V_j)^T,
 * then B <- w B^-1(Tikhonov, 1e-8) + (1 - w) I with w = det / (det + max(alpha - det, 0)); variable "LinearCorrectionMatrix" (ndim x ndim) */
sphb200_status sphb200_ck_linear_correction_matrix(sphb200_relation *inner, sphb200_relation *contact_or_null, double alpha);
/* FreeSurfaceIndicationCK<Inner<WithUpdate>> + <Contact> (general_dynamics/surface_indication/surface_indication_ck.hpp:55-170): variables
 * "PositionDivergence", "Indicator", "PreviousSurfaceIndicator" (the reference's int variables are Real-typed 0 / 1 here) */
sphb200_status sphb200_ck_free_surface_indication(sphb200_relation *inner, sphb200_relation *contact_or_null);
/* The CK acoustic sub-step with linear kernel correction - the scheme of tests/tests_sycl/.../test_3d_dambreak_sycl and its 2-D
 * twin (state: Compression, CompressionRate, Displacement; Position is committed once per advection step):
 *   ck_advection_step_setup        AdvectionStepSetup::UpdateKernel (fluid_time_step_ck.h:145-148): Vol = m / rho, Displacement = 0
 *   ck_update_particle_position    UpdateParticlePosition::UpdateKernel (fluid_time_step_ck.h:173-175): Position += Displacement
 *   ck_acoustic_dt                 ReduceDynamicsCK<AcousticTimeStepCK> (fluid_time_step_ck.hpp:31-35,52-56):
 *                                  cfl h / (max_i max(c0 + |v|, sqrt(4 h |F + Fp| / m)) + tiny)
 *   ck_advection_dt                ReduceDynamicsCK<AdvectionTimeStepCK> (fluid_time_step_ck.cpp:23-25, .h:106-108)
 *   ck_acoustic_step_1st_half      InteractionDynamicsCK<AcousticStep1stHalfWithWallRiemannCorrectionCK>::exec(dt)
 *                                  (acoustic_step_1st_half.hpp:66-74,89-111,122-127,157-180); needs LinearCorrectionMatrix
 *   ck_acoustic_step_2nd_half      ...2ndHalf... (acoustic_step_2nd_half.hpp:33-38,53-73,83-89,117-137)
 * wall_or_null: contact relation to Solid walls (their Velocity / Acceleration are the averages the wall model reads,
 * base_material.cpp:57-65; absent = zero; NormalDirection is required by the 2nd half). */
/* Force of the fluid on a structure.  `contact`: a contact relation whose SOURCE is the solid body and whose
 * targets are fluid bodies.  InteractionDynamicsCK<..., FSI::PressureForceOnStructure<AcousticStep2ndHalfType>> and
 * FSI::ViscousForceOnStructure<ViscousForceType> (shared_ck/particle_dynamics/fluid_structure_interaction/
 * force_on_structure.hpp:113-137 and :61-79), including their WithUpdate step ForcePriorCK::UpdateKernel
 * (general_dynamics/force_prior_ck.h:55-59: ForcePrior += F - PreviousF).  Results in the solid's
 * "PressureForceFromFluid" / "ViscousForceFromFluid".  The solid's Velocity / Acceleration are the averages the structure
 * model reads (absent = zero); NormalDirection is required by the pressure force.  corrected != 0: the fluid's
 * LinearCorrectionMatrix multiplies e_ij (LinearCorrectionCK), 0: NoKernelCorrectionCK. */
sphb200_status sphb200_ck_pressure_force_from_fluid(sphb200_relation *contact, int corrected);
sphb200_status sphb200_ck_viscous_force_from_fluid(sphb200_relation *contact, int corrected);
sphb200_status sphb200_ck_advection_step_setup(sphb200_body *body);
sphb200_status sphb200_ck_update_particle_position(sphb200_body *body);
sphb200_status sphb200_ck_acoustic_dt(sphb200_body *body, double cfl, double *dt_out);
sphb200_status sphb200_ck_advection_dt(sphb200_body *body, double U_ref, double cfl, double *dt_out);
sphb200_status sphb200_ck_acoustic_step_1st_half(sphb200_relation *inner, sphb200_relation *wall_or_null, double dt);
sphb200_status sphb200_ck_acoustic_step_2nd_half(sphb200_relation *inner, sphb200_relation *wall_or_null, double dt);

/* ---- total-Lagrangian elastic solid -------------------------------------------------------- */
/* LinearGradientCorrectionMatrix<Inner<>> (general_dynamics/kernel_correction.cpp:8-29) */
sphb200_status sphb200_correction_matrix(sphb200_relation *inner);
/* Dynamics1Level<DecomposedIntegration1stHalf | Integration1stHalfPK2>
 * (solid_dynamics/elastic_dynamics.h:171-207,259-287; elastic_dynamics.cpp:96-118,162-184) */
sphb200_status sphb200_solid_1st_half(sphb200_relation *inner, int decomposed, double dt);
/* Dynamics1Level<solid_dynamics::Integration2ndHalf> (elastic_dynamics.h:325-351, .cpp:186-194) */
sphb200_status sphb200_solid_2nd_half(sphb200_relation *inner, double dt);
/* SimpleDynamics<FixBodyPartConstraint> over ORIGINAL particle ids (general_constraint.h:106-121) */
sphb200_status sphb200_solid_constrain(sphb200_body *b, const int64_t *original_ids, int64_t n);
/* ReduceDynamics<solid_dynamics::AcousticTimeStep> (elastic_dynamics.cpp:12-30), ReduceMin */
sphb200_status sphb200_solid_dt(sphb200_body *b, double cfl, double *dt_out);

/* ---- primitives exposed for the reference's own known-answer tests -------------------------- */
/* exclusive_scan(policy, first, d_first, n, plus) (common/algorithm_primitive.h:209-279; SYCL twin
 * src_sycl/shared/common/algorithm_primitive_sycl.h:96-122).  Host in/out buffers of n uint32;
 * returns the total like the reference's overload. */
sphb200_status sphb200_exclusive_scan_u32(sphb200_ctx *ctx, const uint32_t *in, uint32_t *out, int64_t n, uint64_t *total);

/* The same scan for callers that cannot allocate.  It has NO counterpart in the reference and was added
 * (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  DEVICE pointers:
 * d_in holds n uint32, d_out n + 1 (d_out[i] = sum of d_in[0..i), d_out[n] = the total; in place, d_out == d_in, is not
 * supported); enqueued on the context's stream, no host synchronisation.  The call performs no cudaMalloc / cudaFree
 * and calls no allocating library routine: the scan's scratch (one partial sum per tile of 4096 elements) is a
 * file-scope device array of 16 384 tiles, so n <= 67 108 864 - larger n returns SPHB200_ERR_INVALID with a message
 * naming the limit.  Every GPU that loads the library holds those 64 KB until exit.  All calls on a device share the
 * array: calls that overlap on the GPU (two contexts or streams on one device) race, callers must order them.  Nothing
 * in it has to start at zero (every word is written before it is read), so back-to-back calls need no clearing.  The
 * first call per device resolves the array's device address (cudaGetSymbolAddress); do one call (n = 0 will do)
 * before starting a graph capture. */
sphb200_status sphb200_exclusive_scan_u32_static(sphb200_ctx *ctx, const uint32_t *d_in, uint32_t *d_out, int64_t n);

/* ---- slab decomposition helpers (multi-GPU; no reference equivalent) ------------ */
/* enqueue all further work of this context on the caller's cudaStream_t (e.g. torch's current
 * stream, so that torch.distributed collectives and these kernels are ordered without host syncs) */
sphb200_status sphb200_ctx_set_stream(sphb200_ctx *ctx, void *cuda_stream);
/* change the number of particles (migration in/out, halo rebuild); keeps the first
 * min(n_old, n_new) rows of every variable, zero-fills the rest, resets OriginalID/SortedID to iota
 * and invalidates the cell list.  Device pointers obtained earlier are invalid afterwards. */
sphb200_status sphb200_body_resize(sphb200_body *b, int64_t n_new);
/* this rank owns the cell pages [page_lo, page_hi) along axis 0 (the slowest index of the linear cell
 * index, base_mesh.hxx:68-78, so a slab is a contiguous range of the cell-sorted particles).  After
 * the next cell_list_update the dynamics and reductions visit only the owned particles as i; all
 * particles (owned + halo copies) serve as neighbours j.  page_lo < 0 restores "own everything". */
sphb200_status sphb200_body_set_owned_pages(sphb200_body *b, int page_lo, int page_hi);
/* [begin, end) of the owned particles in the current (cell-sorted) order */
sphb200_status sphb200_body_owned_range(sphb200_body *b, int64_t *begin, int64_t *end);
/* device pointers of the u32 OriginalID / SortedID arrays (valid until the next resize) */
sphb200_status sphb200_ids_device_ptr(sphb200_body *b, void **original_id, void **sorted_id);
/* drop every particle outside [begin, end) of the current order (the halo copies, before migration):
 * rows [begin, end) of every variable move to the front, n becomes end-begin, ids reset to iota, the cell
 * list is invalidated */
sphb200_status sphb200_body_keep_range(sphb200_body *b, int64_t begin, int64_t end);
/* index of the first particle (cell-sorted order) whose cell page along axis 0 is >= page; needs a valid
 * cell list.  page <= 0 gives 0, page >= number of pages gives n */
sphb200_status sphb200_body_page_begin(sphb200_body *b, int page, int64_t *out);
/* out[p] = page_begin(p) for p = 0..n_pages (n_pages + 1 values; out may be NULL to query n_pages): the per-page
 * particle counts a re-cut of the slabs needs, in one strided device-to-host copy */
sphb200_status sphb200_body_page_offsets(sphb200_body *b, int64_t *out, int64_t *n_pages);
/* the packed gather records of the acoustic sub-step (Rec8 = 8 Reals per particle {x,y,z,p|Vol,vx,vy,vz},
 * which = 0: rec_a written by the 1st-half initialization, 1: rec_b written by the 1st-half update).  A
 * slab run overwrites the records of its halo copies with the owner's between the stages. */
sphb200_status sphb200_body_record_ptr(sphb200_body *b, int which, void **out);
/* Re-decomposition of a slab after an advection step, without a cell-list update of the old arrangement:
 * slab_select classifies the rows by the axis-0 cell page of their CURRENT position (old halo copies are dropped; owned
 * rows stay if their page is in [page_lo - 1, page_hi + 1) - as owned rows or as this slab's own fresh halo copy of a
 * migrant - go to the left slab if page <= page_lo, to the right one if page >= page_hi - 1 - as migrants or as that
 * slab's fresh halo) and returns counts3 = {stay, to left, to right}.  The pages are the CURRENT owned pages: call
 * sphb200_slab_recut first to move a cut (load re-balancing; a cut may move by less than the width of the
 * slabs either side, so that every row a slab needs is owned by itself or a direct neighbour);
 * slab_pack gathers the listed variables of the rows going to `side` (1 left, 2 right) into a row-major device
 * buffer [count x sum(width)]; slab_rebuild makes the body {staying rows, rows received from the left, from the
 * right} (every variable compacted in one pass, listed variables filled from the received buffers, the others
 * zero), resets the ids and invalidates the cell list: ONE cell_list_update then sorts the new arrangement. */
sphb200_status sphb200_slab_select(sphb200_body *b, int has_left, int has_right, int64_t *counts3);
/* move the cuts of this slab: the owned pages become [page_lo, page_hi) for the next slab_select / cell_list_update,
 * the owned RANGE stays that of the current arrangement until then (unlike body_set_owned_pages, which re-derives it
 * at once) */
sphb200_status sphb200_slab_recut(sphb200_body *b, int page_lo, int page_hi);
sphb200_status sphb200_slab_pack(sphb200_body *b, int side, const char *const *names, int n_names, void *out);
sphb200_status sphb200_slab_rebuild(sphb200_body *b, const char *const *names, int n_names, const void *from_left, int64_t n_left,
                                    const void *from_right, int64_t n_right);
/* halo pack / unpack: dst[r] = src[idx[r]] and dst[idx[r]] = src[r] for n rows of row_bytes (multiple of 16) bytes,
 * all device pointers, on the context's stream */
sphb200_status sphb200_rows_gather(sphb200_ctx *ctx, const void *src, int row_bytes, const int64_t *idx, int64_t n, void *dst);
sphb200_status sphb200_rows_scatter(sphb200_ctx *ctx, void *dst, int row_bytes, const int64_t *idx, int64_t n, const void *src);
/* Halo records over NVLink peer memory (csrc/peer.cu): the pack kernel stores the rows straight into a landing buffer in
 * the neighbour GPU's memory and publishes `epoch` (a counter that grows with every refresh, the same on both sides) in the
 * neighbour's flag word; the neighbour waits for that epoch on ITS stream (bounded spin on the device, no host) and unpacks.
 *   peer_alloc / peer_free     landing buffers and flag words (zero-filled, exportable allocations)
 *   peer_export / peer_open    64-byte handle of an allocation / the mapping of another process's allocation on this device
 *                              (cudaIpcGetMemHandle / cudaIpcOpenMemHandle; inside one process pass the pointer itself)
 *   rows_push                  peer_landing[r] = src[idx[r]] for n rows, then *peer_flag = epoch (release, system scope)
 *   rows_wait_scatter          wait until *flag >= epoch, then dst[idx[r]] = landing[r]
 *   peer_timeouts              number of waits that gave up after ~5 s (a dead neighbour must not hang the device) */
sphb200_status sphb200_peer_alloc(sphb200_ctx *ctx, int64_t bytes, void **out);
sphb200_status sphb200_peer_free(sphb200_ctx *ctx, void *p);
sphb200_status sphb200_peer_export(sphb200_ctx *ctx, const void *p, unsigned char handle64[64]);
sphb200_status sphb200_peer_open(sphb200_ctx *ctx, const unsigned char handle64[64], void **out);
sphb200_status sphb200_peer_close(sphb200_ctx *ctx, void *p);
sphb200_status sphb200_rows_push(sphb200_ctx *ctx, const void *src, int row_bytes, const int64_t *idx, int64_t n, void *peer_landing,
                                 void *peer_flag, uint64_t epoch);
sphb200_status sphb200_rows_wait_scatter(sphb200_ctx *ctx, void *dst, int row_bytes, const int64_t *idx, int64_t n, const void *landing,
                                         const void *flag, uint64_t epoch);
sphb200_status sphb200_peer_timeouts(sphb200_ctx *ctx, int64_t *out);
/* Dynamics1Level<Integration1stHalf...>::exec split into its two particle loops
 * (particle_dynamics/dynamics_algorithms.h:336-352): initialization (fluid_integration.hpp:50-55), then
 * interaction + update (:58-113), so that a slab run can refresh the halo pressures in between */
sphb200_status sphb200_integration_1st_half_init(sphb200_relation *inner, double dt);
sphb200_status sphb200_integration_1st_half_interact(sphb200_relation *inner, sphb200_relation *wall_or_null, int riemann, double dt);


#ifdef __cplusplus
}
#endif
#endif /* SPHB200_H */
