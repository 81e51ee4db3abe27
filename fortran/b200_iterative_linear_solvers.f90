! b200_ils_t: the eight entries of the reference's solver factory ('CG', 'ICG', 'RICHARDSON', 'RGMRES', 'FGMRES', 'LGMRES', 'LFOM',
! 'MINRES'; iterative_linear_solver.f90:62-84) on the device.  Extends base_iterative_linear_solver_t and implements its six
! deferred TBPs (base_iterative_linear_solver.f90:160-165); registered after FEMPAR_INIT with
!   call the_iterative_linear_solver_creational_methods_dictionary%set('CG', create_b200_cg)   (..._dictionary.f90:97-106)
! solve_body: the device runs the iterations ahead of the host and returns the iteration count, the convergence flag, the
! error-estimate history and the tolerances; the reference's convergence log is then printed with the base class's own
! print_convergence_history_header / _new_line / _footer (base_iterative_linear_solver.f90:584-676), honouring
! ILS_OUTPUT_FREQUENCY and ILS_LUOUT, so that logs diff cleanly against the CPU path.
! SOURCE ONLY (no Fortran compiler in the build image); mirrored by fempar_b200/fempar.py class IterativeLinearSolver.
module b200_iterative_linear_solvers_names
  use types_names
  use vector_names
  use operator_names
  use environment_names
  use base_iterative_linear_solver_names
  use iterative_linear_solver_parameters_names
  use FPL
  use iso_c_binding
  use fempar_b200_bindings
  use b200_scalar_array_names
  use b200_sparse_matrix_names
  use b200_jacobi_preconditioner_names
  implicit none
# include "debug.i90"
  private

  ! FB2_ILS_* of include/fempar_b200.h
  integer(c_int), parameter :: ils_cg = 0, ils_icg = 1, ils_richardson = 2, ils_rgmres = 3, ils_fgmres = 4, ils_lgmres = 5, &
                               ils_lfom = 6, ils_minres = 7

  type, extends(base_iterative_linear_solver_t) :: b200_ils_t
     integer(c_int) :: ils_type = ils_cg
     integer(c_int) :: dkrymax = 1000, orthonorm_strat = 1          ! ICGSRO (iterative_solver_parameters.f90:182-183)
     real(c_double) :: relaxation = 1.0_c_double
     real(c_double), allocatable :: history(:), history_extra(:)
   contains
     procedure :: allocate_workspace            => b200_ils_allocate_workspace
     procedure :: free_workspace                => b200_ils_free_workspace
     procedure :: set_parameters_from_pl        => b200_ils_set_parameters_from_pl
     procedure :: solve_body                    => b200_ils_solve_body
     procedure :: supports_stopping_criteria    => b200_ils_supports_stopping_criteria
     procedure :: get_default_stopping_criteria => b200_ils_get_default_stopping_criteria
  end type b200_ils_t

  public :: b200_ils_t, create_b200_cg, create_b200_icg, create_b200_richardson, create_b200_rgmres, create_b200_fgmres, &
            create_b200_lgmres, create_b200_lfom, create_b200_minres, b200_register_iterative_linear_solvers

contains

  subroutine b200_register_iterative_linear_solvers()                  ! call once after FEMPAR_INIT
    call the_iterative_linear_solver_creational_methods_dictionary%set(cg_name        , create_b200_cg)
    call the_iterative_linear_solver_creational_methods_dictionary%set(icg_name       , create_b200_icg)
    call the_iterative_linear_solver_creational_methods_dictionary%set(richardson_name, create_b200_richardson)
    call the_iterative_linear_solver_creational_methods_dictionary%set(rgmres_name    , create_b200_rgmres)
    call the_iterative_linear_solver_creational_methods_dictionary%set(fgmres_name    , create_b200_fgmres)
    call the_iterative_linear_solver_creational_methods_dictionary%set(lgmres_name    , create_b200_lgmres)
    call the_iterative_linear_solver_creational_methods_dictionary%set(lfom_name      , create_b200_lfom)
    call the_iterative_linear_solver_creational_methods_dictionary%set(minres_name    , create_b200_minres)
  end subroutine b200_register_iterative_linear_solvers

  subroutine create_generic(environment, solver, ils_type, name)
    class(environment_t)                          , intent(in)    :: environment
    class(base_iterative_linear_solver_t), pointer, intent(inout) :: solver
    integer(c_int)                                , intent(in)    :: ils_type
    character(*)                                  , intent(in)    :: name
    type(b200_ils_t), pointer :: s
    allocate(s)
    call s%set_environment(environment)
    call s%set_name(name)
    call s%set_defaults()
    s%ils_type = ils_type
    call s%set_stopping_criteria(s%get_default_stopping_criteria())
    call s%set_state(start)
    solver => s
  end subroutine create_generic

  subroutine create_b200_cg(environment, base_iterative_linear_solver)
    class(environment_t)                          , intent(in)    :: environment
    class(base_iterative_linear_solver_t), pointer, intent(inout) :: base_iterative_linear_solver
    call create_generic(environment, base_iterative_linear_solver, ils_cg, cg_name)
  end subroutine create_b200_cg
  subroutine create_b200_icg(environment, base_iterative_linear_solver)
    class(environment_t)                          , intent(in)    :: environment
    class(base_iterative_linear_solver_t), pointer, intent(inout) :: base_iterative_linear_solver
    call create_generic(environment, base_iterative_linear_solver, ils_icg, icg_name)
  end subroutine create_b200_icg
  subroutine create_b200_richardson(environment, base_iterative_linear_solver)
    class(environment_t)                          , intent(in)    :: environment
    class(base_iterative_linear_solver_t), pointer, intent(inout) :: base_iterative_linear_solver
    call create_generic(environment, base_iterative_linear_solver, ils_richardson, richardson_name)
  end subroutine create_b200_richardson
  subroutine create_b200_rgmres(environment, base_iterative_linear_solver)
    class(environment_t)                          , intent(in)    :: environment
    class(base_iterative_linear_solver_t), pointer, intent(inout) :: base_iterative_linear_solver
    call create_generic(environment, base_iterative_linear_solver, ils_rgmres, rgmres_name)
  end subroutine create_b200_rgmres
  subroutine create_b200_fgmres(environment, base_iterative_linear_solver)
    class(environment_t)                          , intent(in)    :: environment
    class(base_iterative_linear_solver_t), pointer, intent(inout) :: base_iterative_linear_solver
    call create_generic(environment, base_iterative_linear_solver, ils_fgmres, fgmres_name)
  end subroutine create_b200_fgmres
  subroutine create_b200_lgmres(environment, base_iterative_linear_solver)
    class(environment_t)                          , intent(in)    :: environment
    class(base_iterative_linear_solver_t), pointer, intent(inout) :: base_iterative_linear_solver
    call create_generic(environment, base_iterative_linear_solver, ils_lgmres, lgmres_name)
  end subroutine create_b200_lgmres
  subroutine create_b200_lfom(environment, base_iterative_linear_solver)
    class(environment_t)                          , intent(in)    :: environment
    class(base_iterative_linear_solver_t), pointer, intent(inout) :: base_iterative_linear_solver
    call create_generic(environment, base_iterative_linear_solver, ils_lfom, lfom_name)
  end subroutine create_b200_lfom
  subroutine create_b200_minres(environment, base_iterative_linear_solver)
    class(environment_t)                          , intent(in)    :: environment
    class(base_iterative_linear_solver_t), pointer, intent(inout) :: base_iterative_linear_solver
    call create_generic(environment, base_iterative_linear_solver, ils_minres, minres_name)
  end subroutine create_b200_minres

  subroutine b200_ils_allocate_workspace(this)          ! the Krylov work vectors live with the operator on the device
    class(b200_ils_t), intent(inout) :: this
    if (allocated(this%history)) deallocate(this%history, this%history_extra)
    allocate(this%history(this%get_max_num_iterations()), this%history_extra(this%get_max_num_iterations()))
  end subroutine b200_ils_allocate_workspace

  subroutine b200_ils_free_workspace(this)
    class(b200_ils_t), intent(inout) :: this
    if (allocated(this%history)) deallocate(this%history, this%history_extra)
  end subroutine b200_ils_free_workspace

  subroutine b200_ils_set_parameters_from_pl(this, parameter_list)
    class(b200_ils_t)    , intent(inout) :: this
    type(ParameterList_t), intent(in)    :: parameter_list
    integer(ip) :: FPLError, dkrymax
    character(len=:), allocatable :: strat
    real(rp) :: relaxation
    call this%base_iterative_linear_solver_set_parameters_from_pl(parameter_list)   ! ILS_RTOL, ILS_ATOL, ILS_STOPPING_CRITERIUM,
                                                                                     ! ILS_OUTPUT_FREQUENCY, ILS_MAX_NUM_ITERATIONS,
                                                                                     ! ILS_TRACK_CONVERGENCE_HISTORY, ILS_LUOUT
    if (parameter_list%isPresent(ils_max_dim_krylov_basis_key)) then
       FPLError = parameter_list%Get(Key=ils_max_dim_krylov_basis_key, Value=dkrymax); assert(FPLError == 0)
       this%dkrymax = dkrymax
    end if
    if (parameter_list%isPresent(ils_orthonorm_strategy_key)) then
       FPLError = parameter_list%GetAsString(Key=ils_orthonorm_strategy_key, String=strat); assert(FPLError == 0)
       this%orthonorm_strat = merge(0_c_int, 1_c_int, strat == orthonorm_strat_mgsro)
    end if
    if (parameter_list%isPresent(ils_relaxation_key)) then
       FPLError = parameter_list%Get(Key=ils_relaxation_key, Value=relaxation); assert(FPLError == 0)
       this%relaxation = relaxation
    end if
  end subroutine b200_ils_set_parameters_from_pl

  integer(c_int) function stop_code(stopping_criteria)             ! FB2_STOP_* 0..9 (iterative_solver_parameters.f90:37-46)
    character(*), intent(in) :: stopping_criteria
    select case (stopping_criteria)
    case (res_res);                    stop_code = 0
    case (res_rhs);                    stop_code = 1
    case (res_nrmgiven_rhs_nrmgiven);  stop_code = 2
    case (res_nrmgiven_res_nrmgiven);  stop_code = 3
    case (delta_rhs);                  stop_code = 4
    case (delta_delta);                stop_code = 5
    case (delta_rhs_and_res_res);      stop_code = 6
    case (delta_rhs_and_res_rhs);      stop_code = 7
    case (delta_delta_and_res_res);    stop_code = 8
    case (delta_delta_and_res_rhs);    stop_code = 9
    case default;                      stop_code = -1
    end select
  end function stop_code

  subroutine b200_ils_solve_body(this, b, x)
    class(b200_ils_t), intent(inout) :: this
    class(vector_t)  , intent(in)    :: b
    class(vector_t)  , intent(inout) :: x
    class(operator_t), pointer :: A, M
    class(vector_t)  , pointer :: initial_solution
    type(fb2_ils_params_t)     :: prm
    type(fb2_ils_ext_params_t) :: ext
    integer(c_int) :: its(1), conv(1)
    real(c_double) :: err(1), cd(4)
    type(c_ptr)    :: dinv, hist
    integer(ip)    :: i, luout
    integer(ip), pointer :: num_iterations
    logical    , pointer :: did_converge
    real(rp)   , pointer :: rhs_test, rhs_extra, est, est_extra

    A => this%get_A(); M => this%get_M(); initial_solution => this%get_initial_solution()
    luout = this%get_luout()
    prm%rtol = this%get_rtol(); prm%atol = this%get_atol()
    prm%max_num_iterations = this%get_max_num_iterations()
    prm%stopping_criterium = stop_code(this%get_stopping_criteria())
    ! the per-iteration lines of the log need the history even when the caller does not track it
    prm%track_convergence_history = merge(1_c_int, 0_c_int, this%get_track_convergence_history() .or. this%get_output_frequency() /= 0)
    prm%check_every = 8
    ext%dkrymax = this%dkrymax; ext%orthonorm_strat = this%orthonorm_strat; ext%relaxation = this%relaxation
    hist = c_null_ptr
    if (prm%track_convergence_history == 1) hist = c_loc(this%history)

    select type (A)
    class is (b200_sparse_matrix_t)
       dinv = c_null_ptr                                ! identity preconditioner unless M is the device Jacobi
       select type (M)
       class is (b200_jacobi_preconditioner_t)
          if (.not. M%numeric) call M%numerical_setup()
          dinv = M%inverse_diagonal%handle
       end select
       ! base_iterative_linear_solver_t%set_initial_solution keeps its own copy (:340-360); zero by default (:281-282): NULL
       ! lets the device skip the initial SpMV
       select type (initial_solution)
       class is (b200_scalar_array_t)
          call b200_check(fb2_csr_set_initial_solution(A%handle, initial_solution%handle))
       class default
          call b200_check(fb2_csr_set_initial_solution(A%handle, c_null_ptr))
       end select
       select type (b)
       class is (b200_scalar_array_t)
          select type (x)
          class is (b200_scalar_array_t)
             if (this%ils_type == ils_cg) then
                call b200_check(fb2_pcg_solve(A%handle, dinv, b%handle, x%handle, prm, its, conv, err, hist))
             else
                call b200_check(fb2_ils_solve(A%handle, dinv, b%handle, x%handle, this%ils_type, prm, ext, its, conv, err, hist))
             end if
          class default
             check(.false.)
          end select
       class default
          check(.false.)
       end select
       call b200_check(fb2_csr_get_convergence_data(A%handle, cd, merge(c_loc(this%history_extra), c_null_ptr, &
            prm%track_convergence_history == 1 .and. prm%stopping_criterium >= 6 .and. this%ils_type <= ils_icg)))
    class default
       check(.false.)
    end select

    ! hand the results to the base class and print its log
    num_iterations => this%get_pointer_num_iterations(); did_converge => this%get_pointer_did_converge()
    rhs_test  => this%get_pointer_rhs_convergence_test();            rhs_extra => this%get_pointer_rhs_extra_convergence_test()
    est       => this%get_pointer_error_estimate_convergence_test(); est_extra => this%get_pointer_error_estimate_extra_convergence_test()
    rhs_test = cd(1); rhs_extra = cd(2)
    call this%print_convergence_history_header(luout)
    if (prm%track_convergence_history == 1) then
       did_converge = .false.
       do i = 1, its(1)
          num_iterations = i; est = this%history(i)
          if (prm%stopping_criterium >= 6) est_extra = this%history_extra(i)
          if (i == its(1)) did_converge = (conv(1) == 1)
          call this%print_convergence_history_new_line(luout)      ! prints every output_frequency iterations and the last one
       end do
    end if
    num_iterations = its(1); did_converge = (conv(1) == 1); est = cd(3); est_extra = cd(4)
    call this%print_convergence_history_footer(luout)
  end subroutine b200_ils_solve_body

  function b200_ils_supports_stopping_criteria(this, stopping_criteria)
    class(b200_ils_t), intent(in) :: this
    character(len=*) , intent(in) :: stopping_criteria
    logical :: b200_ils_supports_stopping_criteria
    select case (this%ils_type)
    case (ils_cg, ils_icg)                      ! cg.f90:240-255: all ten
       b200_ils_supports_stopping_criteria = stop_code(stopping_criteria) >= 0
    case (ils_richardson)                       ! richardson.f90:207-214
       b200_ils_supports_stopping_criteria = (stopping_criteria == res_res .or. stopping_criteria == res_rhs)
    case (ils_minres)                           ! minres.f90:764-770
       b200_ils_supports_stopping_criteria = (stopping_criteria == res_res)
    case default                                ! rgmres/fgmres/lgmres/lfom: the four residual criteria
       b200_ils_supports_stopping_criteria = stop_code(stopping_criteria) >= 0 .and. stop_code(stopping_criteria) <= 3
    end select
  end function b200_ils_supports_stopping_criteria

  function b200_ils_get_default_stopping_criteria(this)
    class(b200_ils_t), intent(in) :: this
    character(len=:), allocatable :: b200_ils_get_default_stopping_criteria
    select case (this%ils_type)                 ! iterative_solver_parameters.f90:173-180
    case (ils_rgmres, ils_fgmres, ils_lgmres)
       b200_ils_get_default_stopping_criteria = res_nrmgiven_res_nrmgiven
    case default
       b200_ils_get_default_stopping_criteria = res_res
    end select
  end function b200_ils_get_default_stopping_criteria
end module b200_iterative_linear_solvers_names
