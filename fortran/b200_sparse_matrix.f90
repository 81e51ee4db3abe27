! b200_sparse_matrix_t: the CSR matrix of fe_affine_operator_t on the device.
! Extends matrix_t (Sources/Lib/LinearAlgebra/SparseMatrix/matrix.f90:35-45; operator_t deferred apply / apply_add / is_linear,
! LinearOperator/operator.f90:39-52).  Two ways in:
!   * create_on_space: the CSR graph is built on the device from a device-numbered structured space (fb2_csr_create), the
!     values come from fb2_integrate_assemble (b200_sparse_assembler.f90);
!   * create_from_csr: a csr_sparse_matrix_t the reference assembled itself -- what sparse_matrix%convert ->
!     move_from_coo_body yields (sparse_matrix.f90:783-833, csr_sparse_matrix.f90:416-479) -- is handed over as it is stored
!     (irp(1) = 1, ja ascending per row, 1-based): unstructured meshes, other reference FEs.
! apply_body = csr matvec (csr_sparse_matrix.f90:545-568) + comm() in par (par_sparse_matrix.f90:739-762).
! SOURCE ONLY (no Fortran compiler in the build image); mirrored by fempar_b200/fempar.py class SparseMatrix.
module b200_sparse_matrix_names
  use types_names
  use memory_guard_names
  use vector_names
  use vector_space_names
  use operator_names
  use matrix_names
  use iso_c_binding
  use fempar_b200_bindings
  use b200_scalar_array_names
  implicit none
# include "debug.i90"
  private

  type, extends(matrix_t) :: b200_sparse_matrix_t
     type(c_ptr) :: handle = c_null_ptr       ! fb2_matrix_t
     type(c_ptr) :: space  = c_null_ptr       ! fb2_space_t (borrowed)
     integer(ip) :: sign   = SPARSE_MATRIX_SIGN_POSITIVE_DEFINITE
     logical     :: symmetric_storage = .false.
   contains
     procedure :: create_on_space   => b200_sparse_matrix_create_on_space
     procedure :: create_from_csr   => b200_sparse_matrix_create_from_csr
     procedure :: allocate          => b200_sparse_matrix_allocate
     procedure :: init              => b200_sparse_matrix_init
     procedure :: scal              => b200_sparse_matrix_scal
     procedure :: add               => b200_sparse_matrix_add
     procedure :: copy              => b200_sparse_matrix_copy
     procedure :: free_in_stages    => b200_sparse_matrix_free_in_stages
     procedure :: create_iterator   => b200_sparse_matrix_create_iterator
     procedure :: extract_diagonal  => b200_sparse_matrix_extract_diagonal
     procedure :: get_sign          => b200_sparse_matrix_get_sign
     procedure :: apply             => b200_sparse_matrix_apply
     procedure :: apply_add         => b200_sparse_matrix_apply_add
     procedure :: is_linear         => b200_sparse_matrix_is_linear
     procedure :: get_nnz           => b200_sparse_matrix_get_nnz
     procedure :: get_csr_to_host   => b200_sparse_matrix_get_csr_to_host
  end type b200_sparse_matrix_t

  public :: b200_sparse_matrix_t

contains

  subroutine b200_sparse_matrix_create_on_space(this, space, symmetric_storage, sign)
    class(b200_sparse_matrix_t), intent(inout) :: this
    type(c_ptr)                , intent(in)    :: space
    logical                    , intent(in)    :: symmetric_storage
    integer(ip)                , intent(in)    :: sign
    call this%free_in_stages(free_clean)
    this%space = space; this%symmetric_storage = symmetric_storage; this%sign = sign
    call b200_check(fb2_csr_create(this%handle, space, merge(1_c_int, 0_c_int, symmetric_storage)))
  end subroutine b200_sparse_matrix_create_on_space

  subroutine b200_sparse_matrix_create_from_csr(this, space, num_rows, irp, ja, val, symmetric_storage, sign)
    class(b200_sparse_matrix_t), intent(inout) :: this
    type(c_ptr)                , intent(in)    :: space      ! c_null_ptr: a bare space of num_rows DOFs is created inside
    integer(ip)                , intent(in)    :: num_rows
    integer(ip)                , intent(in)    :: irp(num_rows+1), ja(*)
    real(rp), target           , intent(in)    :: val(*)
    logical                    , intent(in)    :: symmetric_storage
    integer(ip)                , intent(in)    :: sign
    call this%free_in_stages(free_clean)
    this%space = space; this%symmetric_storage = symmetric_storage; this%sign = sign
    call b200_check(fb2_csr_create_from_host(this%handle, space, int(num_rows, c_int64_t), irp, ja, c_loc(val), &
         merge(1_c_int, 0_c_int, symmetric_storage)))
  end subroutine b200_sparse_matrix_create_from_csr

  subroutine b200_sparse_matrix_allocate(this)
    class(b200_sparse_matrix_t), intent(inout) :: this
    assert(c_associated(this%handle))
  end subroutine b200_sparse_matrix_allocate

  subroutine b200_sparse_matrix_init(this, alpha)
    class(b200_sparse_matrix_t), intent(inout) :: this
    real(rp)                   , intent(in)    :: alpha
    call b200_check(fb2_csr_initialize_values(this%handle, real(alpha, c_double)))
  end subroutine b200_sparse_matrix_init

  ! scal / add / copy / create_iterator: matrix algebra and entry iterators are not on the hot path (the reference uses them
  ! in time steppers and BDDC set-up); they go through the host copy of the values (fb2_csr_get_values / fb2_csr_set_values)
  subroutine b200_sparse_matrix_scal(this, alpha)
    class(b200_sparse_matrix_t), intent(inout) :: this
    real(rp)                   , intent(in)    :: alpha
    real(rp), allocatable, target :: v(:)
    integer(c_int64_t) :: n(1), nnz(1)
    call b200_check(fb2_csr_get_sizes(this%handle, n, nnz))
    allocate(v(nnz(1)))
    call b200_check(fb2_csr_get_values(this%handle, v)); v = alpha * v
    call b200_check(fb2_csr_set_values(this%handle, v))
  end subroutine b200_sparse_matrix_scal

  subroutine b200_sparse_matrix_add(this, alpha, op1, beta, op2)
    class(b200_sparse_matrix_t), intent(inout) :: this
    real(rp)                   , intent(in)    :: alpha
    class(matrix_t)            , intent(in)    :: op1
    real(rp)                   , intent(in)    :: beta
    class(matrix_t)            , intent(in)    :: op2
    real(rp), allocatable :: v1(:), v2(:)
    integer(c_int64_t) :: n(1), nnz(1)
    select type(op1)
    class is (b200_sparse_matrix_t)
       select type(op2)
       class is (b200_sparse_matrix_t)      ! same graph (same space, same storage): values combine entry by entry
          call b200_check(fb2_csr_get_sizes(this%handle, n, nnz))
          allocate(v1(nnz(1)), v2(nnz(1)))
          call b200_check(fb2_csr_get_values(op1%handle, v1)); call b200_check(fb2_csr_get_values(op2%handle, v2))
          v1 = alpha * v1 + beta * v2
          call b200_check(fb2_csr_set_values(this%handle, v1))
       class default
          check(.false.)
       end select
    class default
       check(.false.)
    end select
  end subroutine b200_sparse_matrix_add

  subroutine b200_sparse_matrix_copy(this, op)
    class(b200_sparse_matrix_t), intent(inout) :: this
    class(matrix_t)            , intent(in)    :: op
    real(rp), allocatable :: v(:)
    integer(c_int64_t) :: n(1), nnz(1)
    select type(op)
    class is (b200_sparse_matrix_t)
       if (.not. c_associated(this%handle)) call this%create_on_space(op%space, op%symmetric_storage, op%sign)
       call b200_check(fb2_csr_get_sizes(op%handle, n, nnz))
       allocate(v(nnz(1)))
       call b200_check(fb2_csr_get_values(op%handle, v)); call b200_check(fb2_csr_set_values(this%handle, v))
    class default
       check(.false.)
    end select
  end subroutine b200_sparse_matrix_copy

  subroutine b200_sparse_matrix_create_iterator(this, iblock, jblock, iterator)
    class(b200_sparse_matrix_t)        , intent(in)    :: this
    integer(ip)                        , intent(in)    :: iblock, jblock
    class(matrix_iterator_t), allocatable, intent(inout) :: iterator
    check(.false.)     ! entry iterators walk host memory: export with get_csr_to_host and iterate a csr_sparse_matrix_t
  end subroutine b200_sparse_matrix_create_iterator

  subroutine b200_sparse_matrix_extract_diagonal(this, diagonal)          ! csr_sparse_matrix.f90:4231-4256
    class(b200_sparse_matrix_t), intent(in)    :: this
    real(rp), allocatable      , intent(inout) :: diagonal(:)
    type(b200_scalar_array_t) :: d
    integer(c_int64_t) :: n(1), nnz(1)
    call b200_check(fb2_csr_get_sizes(this%handle, n, nnz))
    if (allocated(diagonal)) deallocate(diagonal)
    allocate(diagonal(n(1)))
    call b200_check(fb2_vec_create_n(d%handle, n(1))); d%n = n(1)
    call b200_check(fb2_extract_diagonal(this%handle, d%handle))
    call d%get_entries_to_host(diagonal)
    call d%free_in_stages(free_clean)
  end subroutine b200_sparse_matrix_extract_diagonal

  function b200_sparse_matrix_get_sign(this) result(sign)
    class(b200_sparse_matrix_t), intent(in) :: this
    integer(ip) :: sign
    sign = this%sign
  end function b200_sparse_matrix_get_sign

  subroutine b200_sparse_matrix_apply(this, x, y)                         ! y <- A x (+ comm in par)
    class(b200_sparse_matrix_t), intent(inout) :: this
    class(vector_t)            , intent(in)    :: x
    class(vector_t)            , intent(inout) :: y
    call x%GuardTemp()
    select type(x)
    class is (b200_scalar_array_t)
       select type(y)
       class is (b200_scalar_array_t)
          call b200_check(fb2_spmv(this%handle, x%handle, y%handle))
       class default
          check(.false.)
       end select
    class default
       check(.false.)
    end select
    call x%CleanTemp()
  end subroutine b200_sparse_matrix_apply

  subroutine b200_sparse_matrix_apply_add(this, x, y)                     ! y <- y + A x
    class(b200_sparse_matrix_t), intent(inout) :: this
    class(vector_t)            , intent(in)    :: x
    class(vector_t)            , intent(inout) :: y
    type(b200_scalar_array_t) :: w
    call x%GuardTemp()
    call w%clone(y)
    call this%apply(x, w)
    call y%axpby(1.0_rp, w, 1.0_rp)
    call w%free_in_stages(free_clean)
    call x%CleanTemp()
  end subroutine b200_sparse_matrix_apply_add

  function b200_sparse_matrix_is_linear(this) result(is_linear)
    class(b200_sparse_matrix_t), intent(in) :: this
    logical :: is_linear
    is_linear = .true.
  end function b200_sparse_matrix_is_linear

  function b200_sparse_matrix_get_nnz(this) result(nnz)
    class(b200_sparse_matrix_t), intent(in) :: this
    integer(c_int64_t) :: nnz, n(1), z(1)
    call b200_check(fb2_csr_get_sizes(this%handle, n, z)); nnz = z(1)
  end function b200_sparse_matrix_get_nnz

  ! export in the reference's conventions: irp(1) = 1 (int64: config 2 has 8.48e9 non-zeros > integer(ip)), ja 1-based ascending
  subroutine b200_sparse_matrix_get_csr_to_host(this, irp, ja, val)
    class(b200_sparse_matrix_t), intent(in)    :: this
    integer(c_int64_t), target , intent(inout) :: irp(:)
    integer(ip)       , target , intent(inout) :: ja(:)
    real(rp)                   , intent(inout) :: val(:)
    call b200_check(fb2_csr_get_pattern(this%handle, c_loc(irp), c_loc(ja)))
    call b200_check(fb2_csr_get_values(this%handle, val))
  end subroutine b200_sparse_matrix_get_csr_to_host

  subroutine b200_sparse_matrix_free_in_stages(this, action)
    class(b200_sparse_matrix_t), intent(inout) :: this
    integer(ip)                , intent(in)    :: action
    if (action == free_clean .and. c_associated(this%handle)) then
       call b200_check(fb2_csr_destroy(this%handle))      ! also releases the solver workspace kept with the operator
       this%handle = c_null_ptr
    end if
  end subroutine b200_sparse_matrix_free_in_stages
end module b200_sparse_matrix_names
