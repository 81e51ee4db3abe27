! b200_jacobi_preconditioner_t: jacobi_preconditioner_t (Sources/Lib/LinearAlgebra/Preconditioner/jacobi_preconditioner.f90:
! 136-202) with the inverse diagonal on the device.  numerical_setup (:356-412) = extract_diagonal
! (csr_sparse_matrix.f90:4231-4256) + comm() of the diagonal in par + entrywise_invert; apply = entrywise product.
! SOURCE ONLY (no Fortran compiler in the build image); mirrored by fempar_b200/fempar.py class JacobiPreconditioner.
module b200_jacobi_preconditioner_names
  use types_names
  use memory_guard_names
  use vector_names
  use operator_names
  use iso_c_binding
  use fempar_b200_bindings
  use b200_scalar_array_names
  use b200_sparse_matrix_names
  implicit none
# include "debug.i90"
  private

  type, extends(operator_t) :: b200_jacobi_preconditioner_t
     type(b200_sparse_matrix_t), pointer :: matrix => null()
     type(b200_scalar_array_t)           :: inverse_diagonal
     logical                             :: numeric = .false.
   contains
     procedure :: create          => b200_jacobi_create
     procedure :: symbolic_setup  => b200_jacobi_symbolic_setup
     procedure :: numerical_setup => b200_jacobi_numerical_setup
     procedure :: update_matrix   => b200_jacobi_update_matrix
     procedure :: apply           => b200_jacobi_apply
     procedure :: apply_add       => b200_jacobi_apply_add
     procedure :: is_linear       => b200_jacobi_is_linear
     procedure :: free            => b200_jacobi_free
  end type b200_jacobi_preconditioner_t

  public :: b200_jacobi_preconditioner_t

contains

  subroutine b200_jacobi_create(this, matrix)
    class(b200_jacobi_preconditioner_t), intent(inout) :: this
    type(b200_sparse_matrix_t), target , intent(in)    :: matrix
    call this%free()
    this%matrix => matrix
  end subroutine b200_jacobi_create

  subroutine b200_jacobi_symbolic_setup(this)                          ! create_and_allocate_inverse_diagonal
    class(b200_jacobi_preconditioner_t), intent(inout) :: this
    assert(associated(this%matrix))
    if (.not. c_associated(this%inverse_diagonal%handle)) call this%inverse_diagonal%create_on_space(this%matrix%space)
  end subroutine b200_jacobi_symbolic_setup

  subroutine b200_jacobi_numerical_setup(this)
    class(b200_jacobi_preconditioner_t), intent(inout) :: this
    call this%symbolic_setup()
    call b200_check(fb2_jacobi_setup(this%matrix%handle, this%inverse_diagonal%handle))
    this%numeric = .true.
  end subroutine b200_jacobi_numerical_setup

  subroutine b200_jacobi_update_matrix(this, same_nonzero_pattern)     ! jacobi_preconditioner.f90:200: values changed
    class(b200_jacobi_preconditioner_t), intent(inout) :: this
    logical                            , intent(in)    :: same_nonzero_pattern
    call this%numerical_setup()
  end subroutine b200_jacobi_update_matrix

  subroutine b200_jacobi_apply(this, x, y)                             ! y <- inv(D) x
    class(b200_jacobi_preconditioner_t), intent(inout) :: this
    class(vector_t)                    , intent(in)    :: x
    class(vector_t)                    , intent(inout) :: y
    if (.not. this%numeric) call this%numerical_setup()
    call x%GuardTemp()
    call y%entrywise_product(this%inverse_diagonal, x)
    call x%CleanTemp()
  end subroutine b200_jacobi_apply

  subroutine b200_jacobi_apply_add(this, x, y)                         ! y <- y + inv(D) x
    class(b200_jacobi_preconditioner_t), intent(inout) :: this
    class(vector_t)                    , intent(in)    :: x
    class(vector_t)                    , intent(inout) :: y
    type(b200_scalar_array_t) :: w
    call x%GuardTemp()
    call w%clone(y)
    call this%apply(x, w)
    call y%axpby(1.0_rp, w, 1.0_rp)
    call w%free_in_stages(free_clean)
    call x%CleanTemp()
  end subroutine b200_jacobi_apply_add

  function b200_jacobi_is_linear(this) result(is_linear)
    class(b200_jacobi_preconditioner_t), intent(in) :: this
    logical :: is_linear
    is_linear = .true.
  end function b200_jacobi_is_linear

  subroutine b200_jacobi_free(this)
    class(b200_jacobi_preconditioner_t), intent(inout) :: this
    call this%inverse_diagonal%free_in_stages(free_clean)
    nullify(this%matrix); this%numeric = .false.
  end subroutine b200_jacobi_free
end module b200_jacobi_preconditioner_names
