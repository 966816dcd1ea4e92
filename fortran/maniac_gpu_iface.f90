!> ISO_C_BINDING interface to libmaniacgpu.so (include/maniac_gpu.h).
!> Add this file to the reference's src/ (Makefile:13 picks up src/*.f90) and link with
!>   -L<repo>/maniac_mc_b200 -lmaniacgpu -lcudart -lstdc++
!> COMPILE-UNTESTED: neither the build image nor the GPU box has a Fortran compiler.
module maniac_gpu_iface
    use, intrinsic :: iso_c_binding
    implicit none
    private
    public :: mmc_desc, mmc_init, mmc_destroy, mmc_last_error, mmc_system_energy, mmc_trial_energy, &
              mmc_trial_commit, mmc_sync_to_host, mmc_run_batch, mmc_ewald_setup, mmc_check, &
              mmc_set_chain_params, mmc_run_batch_async, mmc_synchronize, mmc_download_chains, mmc_upload_chains, &
              mmc_shard_step_begin, mmc_shard_step_end, mmc_shard_trial_begin, mmc_shard_trial_end, mmc_trial_energy_swap, mmc_adjust_step_sizes

    integer(c_int), parameter, public :: MMC_CREATION = 1, MMC_DELETION = 2, MMC_TRANSLATION = 3, MMC_ROTATION = 4, MMC_SWAP = 5
    integer(c_int), parameter, public :: MMC_NE = 6

    !> struct mmc_desc (field order and types must match include/maniac_gpu.h)
    type, bind(C) :: mmc_desc
        integer(c_int) :: nres, max_atoms, ntypes, n_chains
        type(c_ptr) :: natoms, is_active, count, capacity
        type(c_ptr) :: charges, types, sigma, epsilon, com, offsets
        real(c_double) :: box_lo(3), box_len(3)
        real(c_double) :: alpha, cutoff, temperature
        type(c_ptr) :: fugacity
        integer(c_int) :: kmax(3)
        real(c_double) :: translation_step, rotation_step, p_translation, p_rotation
        integer(c_int) :: init_sk, deletion_mode, use_table, device
        real(c_double) :: coulomb_xcut
        integer(c_int) :: shard_rank, shard_count
        real(c_double) :: p_swap
        type(c_ptr) :: reservoir_count, reservoir_offsets
        real(c_double) :: tilt(3)
    end type mmc_desc

    interface
        integer(c_int) function mmc_ewald_setup(cutoff, tolerance, box_len, alpha, kmax, nk) bind(C, name="mmc_ewald_setup")
            import :: c_int, c_double
            real(c_double), value :: cutoff, tolerance
            real(c_double), intent(in) :: box_len(3)
            real(c_double), intent(out) :: alpha
            integer(c_int), intent(out) :: kmax(3), nk
        end function
        integer(c_int) function mmc_init(handle, desc) bind(C, name="mmc_init")
            import :: c_int, c_ptr, mmc_desc
            type(c_ptr), intent(out) :: handle
            type(mmc_desc), intent(in) :: desc
        end function
        subroutine mmc_destroy(handle) bind(C, name="mmc_destroy")
            import :: c_ptr
            type(c_ptr), value :: handle
        end subroutine
        type(c_ptr) function mmc_last_error() bind(C, name="mmc_last_error")
            import :: c_ptr
        end function
        integer(c_int) function mmc_system_energy(handle, chain, out) bind(C, name="mmc_system_energy")
            import :: c_int, c_ptr, c_double
            type(c_ptr), value :: handle
            integer(c_int), value :: chain
            real(c_double), intent(out) :: out(6)
        end function
        integer(c_int) function mmc_trial_energy(handle, chain, kind, res, mol, com_new, offs_new, old_e, new_e) &
                bind(C, name="mmc_trial_energy")
            import :: c_int, c_ptr, c_double
            type(c_ptr), value :: handle
            integer(c_int), value :: chain, kind, res, mol
            real(c_double), intent(in) :: com_new(3), offs_new(*)
            real(c_double), intent(out) :: old_e(6), new_e(6)
        end function
        integer(c_int) function mmc_trial_energy_swap(handle, chain, res_old, mol, res_new, com_new, offs_new, old_e, new_e) &
                bind(C, name="mmc_trial_energy_swap")
            import :: c_int, c_ptr, c_double
            type(c_ptr), value :: handle
            integer(c_int), value :: chain, res_old, mol, res_new
            real(c_double), intent(in) :: com_new(3), offs_new(*)
            real(c_double), intent(out) :: old_e(6), new_e(6)
        end function
        integer(c_int) function mmc_trial_commit(handle, chain, accept) bind(C, name="mmc_trial_commit")
            import :: c_int, c_ptr
            type(c_ptr), value :: handle
            integer(c_int), value :: chain, accept
        end function
        !> one strict-mode move in ONE launch: energies + mc_acceptance_probability + the test u <= prob + Accept*/Reject*
        integer(c_int) function mmc_trial_step(handle, chain, kind, res, mol, res_new, com_new, offs_new, u, old_e, new_e, prob, accepted) &
                bind(C, name="mmc_trial_step")
            import :: c_int, c_ptr, c_double
            type(c_ptr), value :: handle
            integer(c_int), value :: chain, kind, res, mol, res_new
            real(c_double), intent(in) :: com_new(3), offs_new(*)
            real(c_double), value :: u
            real(c_double), intent(out) :: old_e(6), new_e(6), prob
            integer(c_int), intent(out) :: accepted
        end function
        !> ComputeSystemEnergy of every chain of the handle (framework terms once per handle); out(6, n_chains)
        integer(c_int) function mmc_system_energy_all(handle, out) bind(C, name="mmc_system_energy_all")
            import :: c_int, c_ptr, c_double
            type(c_ptr), value :: handle
            real(c_double), intent(out) :: out(*)
        end function
        integer(c_int) function mmc_sync_to_host(handle, chain, count, com, offsets, energies) bind(C, name="mmc_sync_to_host")
            import :: c_int, c_ptr, c_double
            type(c_ptr), value :: handle
            integer(c_int), value :: chain
            integer(c_int), intent(out) :: count(*)
            real(c_double), intent(out) :: com(*), offsets(*), energies(6)
        end function
        integer(c_int) function mmc_run_batch(handle, n_steps, seed, tape, tape_stride, records, counters) &
                bind(C, name="mmc_run_batch")
            import :: c_int, c_ptr, c_int64_t
            type(c_ptr), value :: handle
            integer(c_int64_t), value :: n_steps, seed, tape_stride
            type(c_ptr), value :: tape, records, counters
        end function
        integer(c_int) function mmc_set_chain_params(handle, chain, fugacity, temperature) bind(C, name="mmc_set_chain_params")
            import :: c_int, c_ptr, c_double
            type(c_ptr), value :: handle
            integer(c_int), value :: chain
            real(c_double), intent(in) :: fugacity(*)
            real(c_double), value :: temperature
        end function
        !> AdjustMoveStepSizes for every chain from its own counters (once per block, batched mode)
        integer(c_int) function mmc_adjust_step_sizes(handle) bind(C, name="mmc_adjust_step_sizes")
            import :: c_int, c_ptr
            type(c_ptr), value :: handle
        end function
        integer(c_int) function mmc_run_batch_async(handle, n_steps, seed) bind(C, name="mmc_run_batch_async")
            import :: c_int, c_ptr, c_int64_t
            type(c_ptr), value :: handle
            integer(c_int64_t), value :: n_steps, seed
        end function
        integer(c_int) function mmc_synchronize(handle) bind(C, name="mmc_synchronize")
            import :: c_int, c_ptr
            type(c_ptr), value :: handle
        end function
        !> bulk per-block exchange of all chains (fixed-capacity arrays, see include/maniac_gpu.h); null = skip
        integer(c_int) function mmc_upload_chains(handle, count, com, offsets) bind(C, name="mmc_upload_chains")
            import :: c_int, c_ptr
            type(c_ptr), value :: handle, count, com, offsets
        end function
        integer(c_int) function mmc_download_chains(handle, count, com, offsets, energies, counters) &
                bind(C, name="mmc_download_chains")
            import :: c_int, c_ptr
            type(c_ptr), value :: handle, count, com, offsets, energies, counters
        end function
        !> k-sharded mode: dev_sums is a DEVICE buffer (6 doubles per chain) all-reduced by the caller between begin and end
        !> per-block snapshot behind the queued work (slot 0 or 1); returns at once
        integer(c_int) function mmc_snapshot_begin(handle, slot) bind(C, name="mmc_snapshot_begin")
            import :: c_ptr, c_int
            type(c_ptr), value :: handle
            integer(c_int), value :: slot
        end function
        !> waits for that snapshot only; the five c_ptr results point into pinned host memory (c_f_pointer them)
        integer(c_int) function mmc_snapshot_wait(handle, slot, count, com, offsets, energies, counters) &
                bind(C, name="mmc_snapshot_wait")
            import :: c_ptr, c_int
            type(c_ptr), value :: handle
            integer(c_int), value :: slot
            type(c_ptr), intent(out) :: count, com, offsets, energies, counters
        end function
        integer(c_int) function mmc_shard_step_begin(handle, seed, dev_sums) bind(C, name="mmc_shard_step_begin")
            import :: c_int, c_ptr, c_int64_t
            type(c_ptr), value :: handle, dev_sums
            integer(c_int64_t), value :: seed
        end function
        integer(c_int) function mmc_shard_step_end(handle, dev_sums, records) bind(C, name="mmc_shard_step_end")
            import :: c_int, c_ptr
            type(c_ptr), value :: handle, dev_sums, records
        end function
        integer(c_int) function mmc_shard_trial_begin(handle, chain, kind, res, mol, com_new, offs_new, dev_sums) &
                bind(C, name="mmc_shard_trial_begin")
            import :: c_int, c_ptr, c_double
            type(c_ptr), value :: handle, dev_sums
            integer(c_int), value :: chain, kind, res, mol
            real(c_double), intent(in) :: com_new(3), offs_new(*)
        end function
        integer(c_int) function mmc_shard_trial_end(handle, chain, dev_sums, old_e, new_e) bind(C, name="mmc_shard_trial_end")
            import :: c_int, c_ptr, c_double
            type(c_ptr), value :: handle, dev_sums
            integer(c_int), value :: chain
            real(c_double), intent(out) :: old_e(6), new_e(6)
        end function
    end interface

contains

    !> non-zero return code -> AbortRun(mmc_last_error()) (output_utils.f90:535)
    subroutine mmc_check(rc)
        use output_utils, only: AbortRun
        integer(c_int), intent(in) :: rc
        character(kind=c_char), pointer :: cmsg(:)
        character(len=512) :: msg
        integer :: i
        if (rc == 0) return
        call c_f_pointer(mmc_last_error(), cmsg, [512])
        msg = ''
        do i = 1, 512
            if (cmsg(i) == c_null_char) exit
            msg(i:i) = cmsg(i)
        end do
        call AbortRun(trim(msg), 1)
    end subroutine mmc_check

end module maniac_gpu_iface
