! b200_sparse_assembler_t + b200_*_discrete_integration_t: the assembly seam.
!   assembler_t deferred TBPs        Sources/Lib/Integration/FEOperator/assembler.f90:37-116
!   sparse_assembler_t semantics     Sources/Lib/Integration/FEOperator/sparse_assembler.f90:45-206
!   discrete_integration_t           Sources/Lib/Integration/FEOperator/discrete_integration.f90:44-75, called from
!                                    fe_affine_operator_compute (sbm_fe_affine_operator.i90:15-34)
! The device integrates whole operators (FB2_KIND_*), so integrate_galerkin has NO cell loop: it hands the operator kind, the
! analytic-data selector (or host-evaluated source samples at the device's quadrature points) and the material to
! fb2_integrate_assemble.  User integrands that are not one of the device kinds keep the element-wise entry points: the
! assembler stages fe_dofs / elmat / elvec of a batch of cells on the host and flushes them with fb2_assemble_batch.
! SOURCE ONLY (no Fortran compiler in the build image); mirrored by fempar_b200/fempar.py (FeAffineOperator.compute,
! SparseMatrix.assembly).
module b200_sparse_assembler_names
  use types_names
  use memory_names
  use assembler_names
  use matrix_names
  use array_names
  use fe_space_names
  use discrete_integration_names
  use function_names
  use iso_c_binding
  use fempar_b200_bindings
  use b200_scalar_array_names
  use b200_sparse_matrix_names
  implicit none
# include "debug.i90"
  private

  integer(ip), parameter :: batch_cells = 4096

  type, extends(assembler_t) :: b200_sparse_assembler_t
     type(b200_sparse_matrix_t), pointer :: b200_matrix => null()
     type(b200_scalar_array_t) , pointer :: b200_array  => null()
     ! staging of the element-wise entry points (one field, nd DOFs per cell)
     integer(ip) :: nd = 0, staged_mat = 0, staged_vec = 0
     integer(c_int32_t), allocatable :: dofs_mat(:,:), dofs_vec(:,:)
     real(c_double)    , allocatable :: elmats(:,:,:), elvecs(:,:)
   contains
     procedure :: set_device_objects      => b200_sparse_assembler_set_device_objects
     procedure :: assembly_array          => b200_sparse_assembler_assembly_array
     procedure :: assembly_matrix         => b200_sparse_assembler_assembly_matrix
     procedure :: compress_storage_matrix => b200_sparse_assembler_compress_storage_matrix
     procedure :: flush                   => b200_sparse_assembler_flush
  end type b200_sparse_assembler_t

  ! discrete_integration_t for the operator kinds the device integrates itself
  type, extends(discrete_integration_t) :: b200_discrete_integration_t
     integer(c_int) :: kind = 0                 ! FB2_KIND_POISSON / _VECTOR_LAPLACIAN / _ELASTICITY
     integer(c_int) :: func = 1                 ! FB2_FUNC_ZERO / _LINEAR / _SHARP_WAVE: analytic Dirichlet + source data
     real(c_double) :: func_params(5) = [200.0_c_double, 0.7_c_double, -0.05_c_double, -0.05_c_double, -0.05_c_double]
     real(c_double) :: material(2)    = [0.0_c_double, 1.0_c_double]      ! lambda, mu
     class(scalar_function_t), pointer :: source_term => null()           ! any user function: sampled on the host
     type(c_ptr) :: space = c_null_ptr
     ! non-linear operator (fe_operator_t): conductivity k(u) = 1 + beta u^2 and the free DOF values of the evaluation point
     real(c_double) :: beta = 0.0_c_double
     type(b200_scalar_array_t), pointer :: evaluation_point => null()
   contains
     procedure :: set_source_term          => b200_discrete_integration_set_source_term
     procedure :: set_device_evaluation_point => b200_discrete_integration_set_device_evaluation_point
     procedure :: integrate_galerkin       => b200_discrete_integration_integrate_galerkin
     procedure :: integrate_residual       => b200_discrete_integration_integrate_residual
     procedure :: integrate_tangent        => b200_discrete_integration_integrate_tangent
  end type b200_discrete_integration_t

  public :: b200_sparse_assembler_t, b200_discrete_integration_t

contains

  subroutine b200_sparse_assembler_set_device_objects(this, matrix, array)
    class(b200_sparse_assembler_t)    , intent(inout) :: this
    type(b200_sparse_matrix_t), target, intent(in)    :: matrix
    type(b200_scalar_array_t) , target, intent(in)    :: array
    this%b200_matrix => matrix; this%b200_array => array
  end subroutine b200_sparse_assembler_set_device_objects

  ! sparse_assembler_t%assembly_array (sparse_assembler.f90:58-86): array(fe_dofs(i)) += elvec(i) for ids in 1..n
  subroutine b200_sparse_assembler_assembly_array(this, num_fields, field_blocks, num_dofs, fe_dofs, elvec)
    class(b200_sparse_assembler_t), intent(inout) :: this
    integer(ip)                   , intent(in)    :: num_fields
    integer(ip)                   , intent(in)    :: field_blocks(num_fields)
    integer(ip)                   , intent(in)    :: num_dofs(num_fields)
    type(i1p_t)                   , intent(in)    :: fe_dofs(num_fields)
    real(rp)                      , intent(in)    :: elvec(:)
    assert(num_fields == 1)
    if (this%nd == 0) call stage_allocate(this, num_dofs(1))
    if (this%staged_vec == batch_cells) call this%flush()
    this%staged_vec = this%staged_vec + 1
    this%dofs_vec(:, this%staged_vec) = fe_dofs(1)%p(1:this%nd)
    this%elvecs(:, this%staged_vec)   = elvec(1:this%nd)
  end subroutine b200_sparse_assembler_assembly_array

  ! sparse_assembler_t%assembly_matrix (sparse_assembler.f90:88-126) -> coo append filter semantics on the device
  ! (base_sparse_matrix.f90:3469-3513): ids < 1 or > n skipped, ia > ja skipped in symmetric storage
  subroutine b200_sparse_assembler_assembly_matrix(this, num_fields, field_blocks, field_coupling, num_row_dofs, num_col_dofs, &
       fe_dofs_row, fe_dofs_col, elmat)
    class(b200_sparse_assembler_t), intent(inout) :: this
    integer(ip)                   , intent(in)    :: num_fields
    integer(ip)                   , intent(in)    :: field_blocks(num_fields)
    logical                       , intent(in)    :: field_coupling(num_fields,num_fields)
    integer(ip)                   , intent(in)    :: num_row_dofs(num_fields)
    integer(ip)                   , intent(in)    :: num_col_dofs(num_fields)
    type(i1p_t)                   , intent(in)    :: fe_dofs_row(num_fields)
    type(i1p_t)                   , intent(in)    :: fe_dofs_col(num_fields)
    real(rp)                      , intent(in)    :: elmat(:,:)
    assert(num_fields == 1 .and. num_row_dofs(1) == num_col_dofs(1))
    if (this%nd == 0) call stage_allocate(this, num_row_dofs(1))
    if (this%staged_mat == batch_cells) call this%flush()
    this%staged_mat = this%staged_mat + 1
    this%dofs_mat(:, this%staged_mat)  = fe_dofs_row(1)%p(1:this%nd)
    this%elmats(:, :, this%staged_mat) = elmat(1:this%nd, 1:this%nd)       ! column-major elmat(idof, jdof)
  end subroutine b200_sparse_assembler_assembly_matrix

  subroutine stage_allocate(this, nd)
    class(b200_sparse_assembler_t), intent(inout) :: this
    integer(ip)                   , intent(in)    :: nd
    this%nd = nd
    allocate(this%dofs_mat(nd, batch_cells), this%dofs_vec(nd, batch_cells))
    allocate(this%elmats(nd, nd, batch_cells), this%elvecs(nd, batch_cells))
  end subroutine stage_allocate

  subroutine b200_sparse_assembler_flush(this)
    class(b200_sparse_assembler_t), intent(inout) :: this
    if (this%staged_mat > 0) then
       call b200_check(fb2_assemble_batch(this%b200_matrix%handle, c_null_ptr, int(this%staged_mat, c_int32_t), &
            int(this%nd, c_int32_t), this%dofs_mat, c_loc(this%elmats), c_null_ptr, c_null_ptr, 0_c_int64_t))
       this%staged_mat = 0
    end if
    if (this%staged_vec > 0) then
       call b200_check(fb2_assemble_batch(this%b200_matrix%handle, this%b200_array%handle, int(this%staged_vec, c_int32_t), &
            int(this%nd, c_int32_t), this%dofs_vec, c_null_ptr, c_loc(this%elvecs), c_null_ptr, 0_c_int64_t))
       this%staged_vec = 0
    end if
  end subroutine b200_sparse_assembler_flush

  ! compress_storage: the CSR graph exists from creation on; what is left is to flush staged cells and to wait for the device
  subroutine b200_sparse_assembler_compress_storage_matrix(this, sparse_matrix_storage_format)
    class(b200_sparse_assembler_t), intent(inout) :: this
    character(*)                  , intent(in)    :: sparse_matrix_storage_format
    assert(sparse_matrix_storage_format == csr_format)
    call this%flush()
    call b200_check(fb2_device_synchronize())
  end subroutine b200_sparse_assembler_compress_storage_matrix

  subroutine b200_discrete_integration_set_source_term(this, source_term)
    class(b200_discrete_integration_t)  , intent(inout) :: this
    class(scalar_function_t), target    , intent(in)    :: source_term
    this%source_term => source_term
  end subroutine b200_discrete_integration_set_source_term

  ! integrate_galerkin(fe_space, assembler): the whole cell loop of the user plugin
  ! (tutorial_01_..._modules.f90:71-137, par_test_poisson_discrete_integration.f90:62-142) as one device call
  subroutine b200_discrete_integration_integrate_galerkin(this, fe_space, assembler)
    class(b200_discrete_integration_t), intent(in)    :: this
    class(serial_fe_space_t)          , intent(inout) :: fe_space
    class(assembler_t)                , intent(inout) :: assembler
    real(c_double), allocatable, target :: xq(:,:,:), fq(:,:)
    type(point_t) , allocatable :: pts(:)
    integer(c_int64_t) :: sizes(8)
    integer(ip) :: c, q, nq, nc
    select type(assembler)
    class is (b200_sparse_assembler_t)
       ! strong Dirichlet values: nodal interpolation of the analytic selector on the device (sbm_serial_fe_space.i90:997-1176)
       call b200_check(fb2_fe_space_interpolate_dirichlet(this%space, this%func, c_loc_or_null(this%func_params)))
       if (associated(this%source_term)) then
          ! a user source term: evaluated on the host at the device's quadrature points, handed over as samples
          call b200_check(fb2_fe_space_get_sizes(this%space, sizes))
          nc = int(sizes(2), ip); nq = nint(real(sizes(3)))            ! one quadrature point per Lagrange node (tensor Gauss rule)
          allocate(xq(3, nq, nc), fq(nq, nc), pts(nq))
          call b200_check(fb2_fe_space_get_quadrature_points(this%space, xq))
          do c = 1, nc
             do q = 1, nq
                call pts(q)%set(1, xq(1,q,c)); call pts(q)%set(2, xq(2,q,c)); call pts(q)%set(3, xq(3,q,c))
             end do
             call this%source_term%get_values_set(pts, fq(:, c))
          end do
          call b200_check(fb2_fe_space_set_source_values(this%space, c_loc(fq)))
       end if
       call b200_check(fb2_integrate_assemble(assembler%b200_matrix%handle, assembler%b200_array%handle, this%kind, this%func, &
            c_loc_or_null(this%func_params), c_loc_or_null(this%material), 0_c_int))
    class default
       check(.false.)
    end select
  contains
    function c_loc_or_null(a) result(p)
      real(c_double), target, intent(in) :: a(:)
      type(c_ptr) :: p
      p = c_loc(a)
    end function c_loc_or_null
  end subroutine b200_discrete_integration_integrate_galerkin

  ! fe_operator_t%set_evaluation_point hands the fe_function of the current point to the discrete integration
  ! (sbm_fe_operator.i90:524-541); on the device it is the scalar array of its free DOF values
  subroutine b200_discrete_integration_set_device_evaluation_point(this, free_dof_values)
    class(b200_discrete_integration_t), intent(inout) :: this
    type(b200_scalar_array_t), target , intent(in)    :: free_dof_values
    this%evaluation_point => free_dof_values
  end subroutine b200_discrete_integration_set_device_evaluation_point

  ! integrate_residual (discrete_integration.f90:44-75; cell loop of test_transient_poisson_discrete_integration.f90:248-319
  ! with k(u_h)): residual_i = sum_K int k(u_h) grad u_h . grad phi_i - f phi_i, into the assembler's array
  subroutine b200_discrete_integration_integrate_residual(this, fe_space, assembler)
    class(b200_discrete_integration_t), intent(in)    :: this
    class(serial_fe_space_t)          , intent(inout) :: fe_space
    class(assembler_t)                , intent(inout) :: assembler
    real(c_double), target :: prm(5)
    assert(associated(this%evaluation_point))
    prm = this%func_params
    select type(assembler)
    class is (b200_sparse_assembler_t)
       call b200_check(fb2_integrate_residual(assembler%b200_matrix%handle, this%evaluation_point%handle, this%func, c_loc(prm), &
            this%beta, assembler%b200_array%handle))
    class default
       check(.false.)
    end select
  end subroutine b200_discrete_integration_integrate_residual

  ! integrate_tangent: tangent_ij = sum_K int k(u_h) grad phi_j . grad phi_i + k'(u_h) phi_j grad u_h . grad phi_i, into the
  ! assembler's matrix (same CSR graph; not symmetric: full storage)
  subroutine b200_discrete_integration_integrate_tangent(this, fe_space, assembler)
    class(b200_discrete_integration_t), intent(in)    :: this
    class(serial_fe_space_t)          , intent(inout) :: fe_space
    class(assembler_t)                , intent(inout) :: assembler
    real(c_double), target :: prm(5)
    assert(associated(this%evaluation_point))
    prm = this%func_params
    select type(assembler)
    class is (b200_sparse_assembler_t)
       call b200_check(fb2_integrate_tangent(assembler%b200_matrix%handle, this%evaluation_point%handle, this%func, c_loc(prm), &
            this%beta))
    class default
       check(.false.)
    end select
  end subroutine b200_discrete_integration_integrate_tangent
end module b200_sparse_assembler_names
