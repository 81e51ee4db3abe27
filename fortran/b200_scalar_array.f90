! b200_scalar_array_t: serial_scalar_array_t / par_scalar_array_t with the entries in the HBM of a B200.
! Extends vector_t (Sources/Lib/LinearAlgebra/Vector/vector.f90:36-52) and forwards its 17 deferred TBPs to the C-ABI of
! libfempar_b200.so (include/fempar_b200.h, interfaces in fempar_b200_bindings.f90).  Reference semantics restated from
! serial_scalar_array.f90:213-419,556-588 and par_scalar_array.f90:303-350,498-533.
! SOURCE ONLY: no Fortran compiler exists in the build image; fempar_b200/fempar.py (class ScalarArray) makes the same calls in
! the same order and is what tests/ exercise.
module b200_scalar_array_names
  use types_names
  use memory_guard_names
  use vector_names
  use iso_c_binding
  use fempar_b200_bindings
  implicit none
# include "debug.i90"
  private

  type, extends(vector_t) :: b200_scalar_array_t
     type(c_ptr) :: handle = c_null_ptr      ! fb2_vector_t
     type(c_ptr) :: space  = c_null_ptr      ! fb2_space_t the vector belongs to (borrowed; gives the par weights / dof_import)
     integer(c_int64_t) :: n = 0
   contains
     procedure :: create_on_space        => b200_scalar_array_create_on_space
     procedure :: allocate               => b200_scalar_array_allocate
     procedure :: dot                    => b200_scalar_array_dot
     procedure :: local_dot              => b200_scalar_array_local_dot
     procedure :: copy                   => b200_scalar_array_copy
     procedure :: init                   => b200_scalar_array_init
     procedure :: scal                   => b200_scalar_array_scal
     procedure :: axpby                  => b200_scalar_array_axpby
     procedure :: nrm2                   => b200_scalar_array_nrm2
     procedure :: clone                  => b200_scalar_array_clone
     procedure :: comm                   => b200_scalar_array_comm
     procedure :: same_vector_space      => b200_scalar_array_same_vector_space
     procedure :: get_num_blocks         => b200_scalar_array_get_num_blocks
     procedure :: extract_subvector      => b200_scalar_array_extract_subvector
     procedure :: insert_subvector       => b200_scalar_array_insert_subvector
     procedure :: entrywise_product      => b200_scalar_array_entrywise_product
     procedure :: entrywise_invert       => b200_scalar_array_entrywise_invert
     procedure :: nullify_non_owned_dofs => b200_scalar_array_nullify_non_owned_dofs
     procedure :: free_in_stages         => b200_scalar_array_free_in_stages
     procedure :: get_entries_to_host    => b200_scalar_array_get_entries_to_host
     procedure :: set_entries_from_host  => b200_scalar_array_set_entries_from_host
  end type b200_scalar_array_t

  public :: b200_scalar_array_t, b200_check

contains

  ! the reference has no error codes: a failing device call aborts like any failed check (Sources/Include/debug.i90)
  subroutine b200_check(ierr)
    integer(c_int), intent(in) :: ierr
    character(kind=c_char), pointer :: msg(:)
    if (ierr /= 0) then
       call c_f_pointer(fb2_last_error_string(), msg, [1024])
       write(0,*) 'libfempar_b200: ', msg(1:index_of_nul(msg)-1)
       check(.false.)
    end if
  contains
    integer function index_of_nul(s)
      character(kind=c_char), intent(in) :: s(:)
      do index_of_nul = 1, size(s)
         if (s(index_of_nul) == c_null_char) return
      end do
    end function index_of_nul
  end subroutine b200_check

  subroutine b200_scalar_array_create_on_space(this, space)
    class(b200_scalar_array_t), intent(inout) :: this
    type(c_ptr)               , intent(in)    :: space
    integer(c_int64_t) :: n(1)
    call this%free_in_stages(free_clean)
    this%space = space
    call b200_check(fb2_vec_create(this%handle, space))
    call b200_check(fb2_vec_size(this%handle, n)); this%n = n(1)
  end subroutine b200_scalar_array_create_on_space

  subroutine b200_scalar_array_allocate(this)          ! entries live on the device from creation on
    class(b200_scalar_array_t), intent(inout) :: this
    assert(c_associated(this%handle))
  end subroutine b200_scalar_array_allocate

  function b200_scalar_array_dot(op1, op2) result(alpha)      ! par: weighted local dot + all-reduce on the device
    class(b200_scalar_array_t), intent(in) :: op1
    class(vector_t)           , intent(in) :: op2
    real(rp) :: alpha
    real(c_double) :: res(1)
    call op1%GuardTemp(); call op2%GuardTemp()
    select type(op2)
    class is (b200_scalar_array_t)
       call b200_check(fb2_vec_dot(op1%handle, op2%handle, res)); alpha = res(1)
    class default
       check(.false.)
    end select
    call op1%CleanTemp(); call op2%CleanTemp()
  end function b200_scalar_array_dot

  function b200_scalar_array_local_dot(op1, op2) result(alpha)
    class(b200_scalar_array_t), intent(in) :: op1
    class(vector_t)           , intent(in) :: op2
    real(rp) :: alpha
    real(c_double) :: res(1)
    call op1%GuardTemp(); call op2%GuardTemp()
    select type(op2)
    class is (b200_scalar_array_t)
       call b200_check(fb2_vec_local_dot(op1%handle, op2%handle, res)); alpha = res(1)
    class default
       check(.false.)
    end select
    call op1%CleanTemp(); call op2%CleanTemp()
  end function b200_scalar_array_local_dot

  subroutine b200_scalar_array_copy(op1, op2)
    class(b200_scalar_array_t), intent(inout) :: op1
    class(vector_t)           , intent(in)    :: op2
    call op2%GuardTemp()
    select type(op2)
    class is (b200_scalar_array_t)
       call b200_check(fb2_vec_copy(op1%handle, op2%handle))
    class default
       check(.false.)
    end select
    call op2%CleanTemp()
  end subroutine b200_scalar_array_copy

  subroutine b200_scalar_array_init(op, alpha)
    class(b200_scalar_array_t), intent(inout) :: op
    real(rp)                  , intent(in)    :: alpha
    call b200_check(fb2_vec_init(op%handle, real(alpha, c_double)))
  end subroutine b200_scalar_array_init

  subroutine b200_scalar_array_scal(op1, alpha, op2)           ! op1 <- alpha * op2
    class(b200_scalar_array_t), intent(inout) :: op1
    real(rp)                  , intent(in)    :: alpha
    class(vector_t)           , intent(in)    :: op2
    call op2%GuardTemp()
    select type(op2)
    class is (b200_scalar_array_t)
       call b200_check(fb2_vec_scal(op1%handle, real(alpha, c_double), op2%handle))
    class default
       check(.false.)
    end select
    call op2%CleanTemp()
  end subroutine b200_scalar_array_scal

  subroutine b200_scalar_array_axpby(op1, alpha, op2, beta)    ! op1 <- alpha*op2 + beta*op1
    class(b200_scalar_array_t), intent(inout) :: op1
    real(rp)                  , intent(in)    :: alpha
    class(vector_t)           , intent(in)    :: op2
    real(rp)                  , intent(in)    :: beta
    call op2%GuardTemp()
    select type(op2)
    class is (b200_scalar_array_t)
       call b200_check(fb2_vec_axpby(op1%handle, real(alpha, c_double), op2%handle, real(beta, c_double)))
    class default
       check(.false.)
    end select
    call op2%CleanTemp()
  end subroutine b200_scalar_array_axpby

  function b200_scalar_array_nrm2(op) result(alpha)
    class(b200_scalar_array_t), intent(in) :: op
    real(rp) :: alpha
    real(c_double) :: res(1)
    call op%GuardTemp()
    call b200_check(fb2_vec_nrm2(op%handle, res)); alpha = res(1)
    call op%CleanTemp()
  end function b200_scalar_array_nrm2

  subroutine b200_scalar_array_clone(op1, op2)
    class(b200_scalar_array_t), target, intent(inout) :: op1
    class(vector_t)           , target, intent(in)    :: op2
    call op2%GuardTemp()
    select type(op2)
    class is (b200_scalar_array_t)
       call op1%create_on_space(op2%space)
    class default
       check(.false.)
    end select
    call op2%CleanTemp()
  end subroutine b200_scalar_array_clone

  subroutine b200_scalar_array_comm(op)                        ! par_scalar_array%comm: 2-stage neighbour exchange over NCCL
    class(b200_scalar_array_t), intent(inout) :: op
    call b200_check(fb2_vec_comm(op%handle))
  end subroutine b200_scalar_array_comm

  function b200_scalar_array_same_vector_space(this, vector)
    class(b200_scalar_array_t), intent(in) :: this
    class(vector_t)           , intent(in) :: vector
    logical :: b200_scalar_array_same_vector_space
    b200_scalar_array_same_vector_space = .false.
    select type(vector)
    class is (b200_scalar_array_t)
       b200_scalar_array_same_vector_space = (this%n == vector%n)
    end select
  end function b200_scalar_array_same_vector_space

  function b200_scalar_array_get_num_blocks(this) result(res)
    class(b200_scalar_array_t), intent(in) :: this
    integer(ip) :: res
    res = 1
  end function b200_scalar_array_get_num_blocks

  subroutine b200_scalar_array_extract_subvector(this, iblock, size_indices, indices, values)    ! explicit D2H
    class(b200_scalar_array_t), intent(in)    :: this
    integer(ip)               , intent(in)    :: iblock
    integer(ip)               , intent(in)    :: size_indices
    integer(ip)               , intent(in)    :: indices(size_indices)
    real(rp)                  , intent(inout) :: values(*)
    assert(iblock == 1)
    call b200_check(fb2_vec_extract_subvector(this%handle, int(size_indices, c_int32_t), indices, values))
  end subroutine b200_scalar_array_extract_subvector

  subroutine b200_scalar_array_insert_subvector(this, iblock, size_indices, indices, values)     ! explicit H2D
    class(b200_scalar_array_t), intent(inout) :: this
    integer(ip)               , intent(in)    :: iblock
    integer(ip)               , intent(in)    :: size_indices
    integer(ip)               , intent(in)    :: indices(size_indices)
    real(rp)                  , intent(in)    :: values(*)
    assert(iblock == 1)
    call b200_check(fb2_vec_insert_subvector(this%handle, int(size_indices, c_int32_t), indices, values))
  end subroutine b200_scalar_array_insert_subvector

  subroutine b200_scalar_array_entrywise_product(op1, op2, op3)
    class(b200_scalar_array_t), intent(inout) :: op1
    class(vector_t)           , intent(in)    :: op2
    class(vector_t)           , intent(in)    :: op3
    call op2%GuardTemp(); call op3%GuardTemp()
    select type(op2)
    class is (b200_scalar_array_t)
       select type(op3)
       class is (b200_scalar_array_t)
          call b200_check(fb2_vec_entrywise_product(op1%handle, op2%handle, op3%handle))
       class default
          check(.false.)
       end select
    class default
       check(.false.)
    end select
    call op2%CleanTemp(); call op3%CleanTemp()
  end subroutine b200_scalar_array_entrywise_product

  subroutine b200_scalar_array_entrywise_invert(op1)
    class(b200_scalar_array_t), intent(inout) :: op1
    call b200_check(fb2_vec_entrywise_invert(op1%handle))
  end subroutine b200_scalar_array_entrywise_invert

  subroutine b200_scalar_array_nullify_non_owned_dofs(this)
    class(b200_scalar_array_t), intent(inout) :: this
    call b200_check(fb2_vec_nullify_non_owned_dofs(this%handle))
  end subroutine b200_scalar_array_nullify_non_owned_dofs

  subroutine b200_scalar_array_free_in_stages(this, action)
    class(b200_scalar_array_t), intent(inout) :: this
    integer(ip)               , intent(in)    :: action
    if (action == free_clean .and. c_associated(this%handle)) then
       call b200_check(fb2_vec_destroy(this%handle))
       this%handle = c_null_ptr; this%n = 0
    end if
  end subroutine b200_scalar_array_free_in_stages

  subroutine b200_scalar_array_get_entries_to_host(this, values)       ! serial_scalar_array%get_entries analogue (D2H copy)
    class(b200_scalar_array_t), intent(in)    :: this
    real(rp), target          , intent(inout) :: values(:)
    assert(size(values) == this%n)
    call b200_check(fb2_vec_get(this%handle, values))
  end subroutine b200_scalar_array_get_entries_to_host

  subroutine b200_scalar_array_set_entries_from_host(this, values)
    class(b200_scalar_array_t), intent(inout) :: this
    real(rp), target          , intent(in)    :: values(:)
    assert(size(values) == this%n)
    call b200_check(fb2_vec_set(this%handle, c_loc(values)))
  end subroutine b200_scalar_array_set_entries_from_host
end module b200_scalar_array_names
