! mcb200.f90 -- ISO_C_BINDING interface to libmcb200.so (include/mcb200.h) for the reference's Fortran driver.
! COMPILE-UNVERIFIED: the build image has no Fortran compiler.  The same argument passing is exercised by the
! ctypes binding (montecarlo_b200/_lib.py) and by the C++ stand-in driver (driver/mcdriver.cpp).
! Where each call goes in MonteCarlo/MonteCarlo.F90 is listed in INTEGRATION.md, section 2.
module mcb200
  use iso_c_binding
  implicit none

  type, bind(C) :: mcb_params            ! include/mcb200.h: mcb_params
    integer(c_int) :: n_mol, n_sites, n_sol, n_chains, device, quirk_31, matrix_mode, lookahead
    real(c_double) :: boxl, cutoff, conv
    type(c_ptr)    :: site_rel, site_tab  ! double[n_sites][3]
  end type

  type, bind(C) :: mcb_mc_params
    real(c_double) :: temperature, dposmax, dposmin, dhoekmax, dhoekmin, padj
    integer(c_int) :: nadj, reserved
    integer(c_int64_t) :: start_prod
  end type

  type, bind(C) :: mcb_chain_stats
    real(c_double) :: toteng, dposmax, dhoekmax, production, sum_e, sum_e2, ratio, last_delta
    integer(c_int64_t) :: step, nsuc, naccept, prod_acc, n_clamped, sitepairs, uniforms_used, n_stat
    integer(c_int64_t) :: exch_attempts, exch_accepted
    real(c_double) :: temperature
  end type

  type, bind(C) :: mcb_global_stats      ! sums over a group of chains, over all ranks (mcb_reduce_stats)
    real(c_double) :: sum_e, sum_e2, production, sum_toteng
    integer(c_int64_t) :: n_stat, prod_acc, nsuc, step, sitepairs, n_clamped, uniforms_used, n_chains
  end type

  integer(c_int), parameter :: MCB_COMM_ID_BYTES = 128

  interface
    integer(c_int) function mcb_create(h, p) bind(C, name="mcb_create")
      import; type(c_ptr), intent(out) :: h; type(mcb_params), intent(in) :: p
    end function
    integer(c_int) function mcb_destroy(h) bind(C, name="mcb_destroy")
      import; type(c_ptr), value :: h
    end function
    type(c_ptr) function mcb_last_error(h) bind(C, name="mcb_last_error")
      import; type(c_ptr), value :: h
    end function
    integer(c_int) function mcb_set_solute(h, chain0, nchains, solute, sol_tab, bcast) bind(C, name="mcb_set_solute")
      import; type(c_ptr), value :: h; integer(c_int), value :: chain0, nchains, bcast
      real(c_double), intent(in) :: solute(*), sol_tab(*)
    end function
    integer(c_int) function mcb_set_chain_state(h, chain0, nchains, com, hoek, bcast) bind(C, name="mcb_set_chain_state")
      import; type(c_ptr), value :: h; integer(c_int), value :: chain0, nchains, bcast
      type(c_ptr), value :: com, hoek            ! c_loc of TYPE(vector) arrays = double[n][3]
    end function
    integer(c_int) function mcb_get_chain_state(h, chain0, nchains, com, hoek) bind(C, name="mcb_get_chain_state")
      import; type(c_ptr), value :: h, com, hoek; integer(c_int), value :: chain0, nchains
    end function
    integer(c_int) function mcb_total_energy(h, chain0, nchains, toteng, esol, pairs) bind(C, name="mcb_total_energy")
      import; type(c_ptr), value :: h, esol, pairs; integer(c_int), value :: chain0, nchains
      real(c_double), intent(out) :: toteng(*)
    end function
    integer(c_int) function mcb_get_matrix(h, chain, ss, energy) bind(C, name="mcb_get_matrix")
      import; type(c_ptr), value :: h; integer(c_int), value :: chain
      real(c_double), intent(out) :: ss(*), energy(*)
    end function
    integer(c_int) function mcb_trial_delta(h, nchains, rsolv, new_com, new_hoek, d_e, e_new) bind(C, name="mcb_trial_delta")
      import; type(c_ptr), value :: h, new_com, new_hoek; integer(c_int), value :: nchains
      integer(c_int), intent(in) :: rsolv(*); real(c_double), intent(out) :: d_e(*), e_new(*)
    end function
    integer(c_int) function mcb_commit(h, nchains, accept) bind(C, name="mcb_commit")
      import; type(c_ptr), value :: h; integer(c_int), value :: nchains
      integer(c_signed_char), intent(in) :: accept(*)
    end function
    integer(c_int) function mcb_set_mc_params(h, chain0, nchains, p, bcast) bind(C, name="mcb_set_mc_params")
      import; type(c_ptr), value :: h; integer(c_int), value :: chain0, nchains, bcast
      type(mcb_mc_params), intent(in) :: p(*)
    end function
    integer(c_int) function mcb_seed_chains(h, chain0, nchains, seeds) bind(C, name="mcb_seed_chains")
      import; type(c_ptr), value :: h; integer(c_int), value :: chain0, nchains
      integer(c_int64_t), intent(in) :: seeds(*)
    end function
    integer(c_int) function mcb_run_steps(h, nsteps, uniforms, cap, consumed, steps_done) bind(C, name="mcb_run_steps")
      import; type(c_ptr), value :: h, uniforms, consumed, steps_done; integer(c_int64_t), value :: nsteps, cap
    end function
    integer(c_int) function mcb_get_stats(h, chain0, nchains, stats) bind(C, name="mcb_get_stats")
      import; type(c_ptr), value :: h; integer(c_int), value :: chain0, nchains
      type(mcb_chain_stats), intent(out) :: stats(*)
    end function
    integer(c_int) function mcb_row_energy(h, chain, i, ss_row, e_i) bind(C, name="mcb_row_energy")
      import; type(c_ptr), value :: h; integer(c_int), value :: chain, i
      real(c_double), intent(out) :: ss_row(*), e_i
    end function
    integer(c_int) function mcb_reset_stats(h, chain0, nchains) bind(C, name="mcb_reset_stats")
      import; type(c_ptr), value :: h; integer(c_int), value :: chain0, nchains
    end function
    integer(c_int) function mcb_set_stream(h, cuda_stream) bind(C, name="mcb_set_stream")
      import; type(c_ptr), value :: h, cuda_stream
    end function
    integer(c_int64_t) function mcb_launch_count(h) bind(C, name="mcb_launch_count")
      import; type(c_ptr), value :: h
    end function
    integer(c_int) function mcb_measure_fp64_peak(device, ms, tflops) bind(C, name="mcb_measure_fp64_peak")
      import; integer(c_int), value :: device; real(c_double), value :: ms; real(c_double), intent(out) :: tflops
    end function
    ! replica exchange between launches: pair_a, pair_b 0-based chain indices; uniforms and accepted may be c_null_ptr
    integer(c_int) function mcb_exchange(h, npairs, pair_a, pair_b, uniforms, accepted) bind(C, name="mcb_exchange")
      import; type(c_ptr), value :: h, uniforms, accepted; integer(c_int), value :: npairs
      integer(c_int), intent(in) :: pair_a(*), pair_b(*)
    end function
    ! end of run: one NCCL all-reduce + one all-gather over the GPUs of the job (one handle per GPU / MPI rank)
    integer(c_int) function mcb_comm_get_unique_id(id) bind(C, name="mcb_comm_get_unique_id")
      import; character(kind=c_char), intent(out) :: id(MCB_COMM_ID_BYTES)
    end function
    integer(c_int) function mcb_comm_init(h, nranks, rank, id, n_total_chains) bind(C, name="mcb_comm_init")
      import; type(c_ptr), value :: h; integer(c_int), value :: nranks, rank
      character(kind=c_char), intent(in) :: id(MCB_COMM_ID_BYTES); integer(c_int64_t), value :: n_total_chains
    end function
    integer(c_int) function mcb_comm_attach(h, nccl_comm, nranks, rank, n_total_chains) bind(C, name="mcb_comm_attach")
      import; type(c_ptr), value :: h, nccl_comm; integer(c_int), value :: nranks, rank; integer(c_int64_t), value :: n_total_chains
    end function
    integer(c_int) function mcb_comm_destroy(h) bind(C, name="mcb_comm_destroy")
      import; type(c_ptr), value :: h
    end function
    integer(c_int) function mcb_reduce_stats(h, chains_per_group, out, n_groups) bind(C, name="mcb_reduce_stats")
      import; type(c_ptr), value :: h; integer(c_int64_t), value :: chains_per_group; integer(c_int), value :: n_groups
      type(mcb_global_stats), intent(out) :: out(*)
    end function
    integer(c_int) function mcb_gather_chains(h, out, n_total_chains) bind(C, name="mcb_gather_chains")
      import; type(c_ptr), value :: h; integer(c_int64_t), value :: n_total_chains
      real(c_double), intent(out) :: out(4, *)     ! (toteng, mean energy, acceptance, production mean) per global chain
    end function
    integer(c_int) function mcb_box_energy_noalloc(p, d_com, d_hoek, solute, sol_tab, d_out, stream, errbuf, errlen) &
        bind(C, name="mcb_box_energy_noalloc")   ! d_com, d_hoek, d_out: DEVICE pointers owned by the caller
      import; type(mcb_params), intent(in) :: p; type(c_ptr), value :: d_com, d_hoek, d_out, stream
      real(c_double), intent(in) :: solute(*), sol_tab(*)
      character(kind=c_char) :: errbuf(*); integer(c_size_t), value :: errlen
    end function
  end interface
end module mcb200
