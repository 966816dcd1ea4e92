!> Glue between MANIAC's module globals (simulation_state.f90) and libmaniacgpu.so.
!>
!> Drop this file and maniac_gpu_iface.f90 into the reference's src/ (Makefile:13 globs src/*.f90).  It provides the
!> routines INTEGRATION.md's call map names:
!>   gpu_attach                 after PrecomputeTable (main.f90:26): packs mmc_desc from primary%, nb%, coeff%, ewald%, input%
!>   gpu_system_energy          body of ComputeSystemEnergy (energy_utils.f90:18)
!>   gpu_trial / gpu_trial_swap ComputeOldEnergy + ComputeNewEnergy of one trial move (monte_carlo_utils.f90:347,275)
!>   gpu_commit                 device side of Accept*/Reject* (monte_carlo_utils.f90:410,485; create_/delete_molecule.f90)
!>   gpu_sync_primary           refreshes primary%mol_com / site_offset / num_residues and `energy` before UpdateFiles /
!>                              PrintStatus (monte_carlo.f90:82-84) in batched mode
!>   gpu_detach                 FinalReport
!> COMPILE-UNTESTED: neither the build image nor the GPU box has a Fortran compiler.  tests/test_abi.py checks by
!> text that the bind(C) type and every bound name match include/maniac_gpu.h and the built library.
module maniac_gpu_glue
    use, intrinsic :: iso_c_binding
    use, intrinsic :: iso_fortran_env, only: real64
    use simulation_state
    use maniac_gpu_iface
    implicit none
    private
    public :: gpu_attach, gpu_system_energy, gpu_trial, gpu_trial_swap, gpu_commit, gpu_sync_primary, gpu_detach, gpu

    type(c_ptr), save :: gpu = c_null_ptr          !< the library handle
    ! packed copies handed to mmc_init (kept alive for the call only; the library copies what it needs)
    integer(c_int), allocatable, target, save :: p_natoms(:), p_active(:), p_count(:), p_cap(:), p_types(:)
    real(c_double), allocatable, target, save :: p_q(:), p_sig(:), p_eps(:), p_com(:), p_off(:), p_fug(:)

contains

    !> Packs (residue, molecule[, atom]) ordered coordinates of the primary box.
    subroutine pack_coordinates()
        integer :: r, m, a, im, ia, nmol, natom
        nmol = 0; natom = 0
        do r = 1, nb%type_residue
            nmol = nmol + primary%num_residues(r)
            natom = natom + primary%num_residues(r) * nb%atom_in_residue(r)
        end do
        if (allocated(p_com)) deallocate(p_com, p_off)
        allocate(p_com(3 * max(nmol, 1)), p_off(3 * max(natom, 1)))
        im = 0; ia = 0
        do r = 1, nb%type_residue
            do m = 1, primary%num_residues(r)
                p_com(3 * im + 1:3 * im + 3) = primary%mol_com(:, r, m)
                im = im + 1
                do a = 1, nb%atom_in_residue(r)
                    p_off(3 * ia + 1:3 * ia + 3) = primary%site_offset(:, r, m, a)
                    ia = ia + 1
                end do
            end do
        end do
    end subroutine pack_coordinates

    subroutine gpu_attach(n_chains, device)
        integer, intent(in) :: n_chains, device
        type(mmc_desc) :: d
        integer :: r, r2, a, a2, ti, tj, nres, maxat, ntypes, i

        nres = nb%type_residue; maxat = nb%max_atom_in_residue; ntypes = primary%num_atomtypes
        allocate(p_natoms(nres), p_active(nres), p_count(nres), p_cap(nres), p_fug(nres))
        allocate(p_q(nres * maxat), p_types(nres * maxat), p_sig(ntypes * ntypes), p_eps(ntypes * ntypes))
        p_q = 0.0_c_double; p_types = 1; p_sig = 0.0_c_double; p_eps = 0.0_c_double
        do r = 1, nres
            p_natoms(r) = nb%atom_in_residue(r)
            p_active(r) = merge(1, 0, input%is_active(r) == 1)
            p_count(r) = primary%num_residues(r)
            p_cap(r) = merge(NB_MAX_MOLECULE, primary%num_residues(r), input%is_active(r) == 1)   ! parameters.f90:8
            p_fug(r) = input%fugacity(r)                          ! already 1/A^3 after prepare_utils.f90:68
            do a = 1, nb%atom_in_residue(r)
                p_q((r - 1) * maxat + a) = primary%atom_charges(r, a)
                p_types((r - 1) * maxat + a) = primary%atom_types(r, a)
            end do
        end do
        ! LJ compacted by ATOM TYPE: the 4-D coeff%sigma/epsilon(res1, res2, atom1, atom2) tables are never copied
        do r = 1, nres
            do r2 = 1, nres
                do a = 1, nb%atom_in_residue(r)
                    do a2 = 1, nb%atom_in_residue(r2)
                        ti = primary%atom_types(r, a); tj = primary%atom_types(r2, a2)
                        p_sig((ti - 1) * ntypes + tj) = coeff%sigma(r, r2, a, a2)
                        p_eps((ti - 1) * ntypes + tj) = coeff%epsilon(r, r2, a, a2)   ! Kelvin (parameters_parser.f90:82)
                    end do
                end do
            end do
        end do
        call pack_coordinates()

        d%nres = nres; d%max_atoms = maxat; d%ntypes = ntypes; d%n_chains = n_chains
        d%natoms = c_loc(p_natoms); d%is_active = c_loc(p_active); d%count = c_loc(p_count); d%capacity = c_loc(p_cap)
        d%charges = c_loc(p_q); d%types = c_loc(p_types); d%sigma = c_loc(p_sig); d%epsilon = c_loc(p_eps)
        d%com = c_loc(p_com); d%offsets = c_loc(p_off); d%fugacity = c_loc(p_fug)
        d%box_lo = primary%bounds(:, 1)
        do i = 1, 3
            d%box_len(i) = primary%matrix(i, i)                   ! lx, ly, lz (the diagonal, also for box%type 3)
        end do
        d%alpha = ewald%alpha; d%cutoff = input%real_space_cutoff; d%temperature = input%temp_K; d%kmax = ewald%kmax
        d%translation_step = input%translation_step; d%rotation_step = input%rotation_step_angle
        d%p_translation = proba%translation; d%p_rotation = proba%rotation; d%p_swap = proba%swap
        d%init_sk = 0; d%deletion_mode = 0; d%use_table = 0       ! reference-compatible defaults (quirks Q1, Q2, Q12)
        d%device = device; d%coulomb_xcut = 0.0_c_double
        d%shard_rank = 0; d%shard_count = 1
        d%reservoir_count = c_null_ptr; d%reservoir_offsets = c_null_ptr   ! strict mode: the host keeps reservoir%
        d%tilt = primary%tilt                                     ! xy, xz, yz; zero for box%type 1 and 2
        call mmc_check(mmc_init(gpu, d))
    end subroutine gpu_attach

    !> ComputeSystemEnergy(box): fills the module-level `energy` exactly as energy_utils.f90:18-35 does
    subroutine gpu_system_energy()
        real(c_double) :: e(6)
        call mmc_check(mmc_system_energy(gpu, 0_c_int, e))
        energy%non_coulomb = e(1); energy%coulomb = e(2); energy%recip_coulomb = e(3)
        energy%ewald_self = e(4); energy%intra_coulomb = e(5); energy%total = e(6)
    end subroutine gpu_system_energy

    subroutine unpack_state(e, s)
        real(c_double), intent(in) :: e(6)
        type(energy_state), intent(out) :: s
        s%non_coulomb = e(1); s%coulomb = e(2); s%recip_coulomb = e(3)
        s%ewald_self = e(4); s%intra_coulomb = e(5); s%total = e(6)
    end subroutine unpack_state

    !> One trial move: `kind` = TYPE_CREATION / TYPE_DELETION / TYPE_TRANSLATION / TYPE_ROTATION (same values as MMC_*).
    !> Call after the proposal has been written to primary%mol_com / site_offset(:, res, mol, :); fills old and new.
    subroutine gpu_trial(kind, res, mol, old_s, new_s)
        integer, intent(in) :: kind, res, mol
        type(energy_state), intent(out) :: old_s, new_s
        real(c_double) :: com(3), off(3 * nb%max_atom_in_residue), eo(6), en(6)
        integer :: a
        com = primary%mol_com(:, res, mol)
        do a = 1, nb%atom_in_residue(res)
            off(3 * a - 2:3 * a) = primary%site_offset(:, res, mol, a)
        end do
        call mmc_check(mmc_trial_energy(gpu, 0_c_int, int(kind, c_int), int(res - 1, c_int), int(mol - 1, c_int), com, off, eo, en))
        call unpack_state(eo, old_s); call unpack_state(en, new_s)
    end subroutine gpu_trial

    !> Swap trial: molecule (res_old, mol) leaves, a molecule of type res_new with the geometry already written to slot
    !> num_residues(res_new) + 1 of the primary box enters (monte_carlo_utils.f90:228-268 gives the acceptance rule).
    subroutine gpu_trial_swap(res_old, mol, res_new, old_s, new_s)
        integer, intent(in) :: res_old, mol, res_new
        type(energy_state), intent(out) :: old_s, new_s
        real(c_double) :: com(3), off(3 * nb%max_atom_in_residue), eo(6), en(6)
        integer :: a, m2
        m2 = primary%num_residues(res_new) + 1
        com = primary%mol_com(:, res_new, m2)
        do a = 1, nb%atom_in_residue(res_new)
            off(3 * a - 2:3 * a) = primary%site_offset(:, res_new, m2, a)
        end do
        call mmc_check(mmc_trial_energy_swap(gpu, 0_c_int, int(res_old - 1, c_int), int(mol - 1, c_int), int(res_new - 1, c_int), &
                                             com, off, eo, en))
        call unpack_state(eo, old_s); call unpack_state(en, new_s)
    end subroutine gpu_trial_swap

    !> Device side of AcceptMove / AcceptCreationMove / AcceptDeletionMove (accepted = .true.) and of the Reject* routines
    !> (.false.: nothing to restore on the device, A(k) is double buffered).
    subroutine gpu_commit(accepted)
        logical, intent(in) :: accepted
        call mmc_check(mmc_trial_commit(gpu, 0_c_int, merge(1_c_int, 0_c_int, accepted)))
    end subroutine gpu_commit

    !> Batched mode, once per block: primary%mol_com / site_offset / num_residues and `energy` from chain `chain` (0-based)
    subroutine gpu_sync_primary(chain)
        integer, intent(in) :: chain
        integer(c_int) :: cnt(nb%type_residue)
        real(c_double), allocatable :: com(:), off(:)
        real(c_double) :: e(6)
        integer :: r, m, a, im, ia, ncap, acap
        ncap = 0; acap = 0
        do r = 1, nb%type_residue
            ncap = ncap + merge(NB_MAX_MOLECULE, primary%num_residues(r), input%is_active(r) == 1)
            acap = acap + merge(NB_MAX_MOLECULE, primary%num_residues(r), input%is_active(r) == 1) * nb%atom_in_residue(r)
        end do
        allocate(com(3 * ncap + 3), off(3 * acap + 3))
        call mmc_check(mmc_sync_to_host(gpu, int(chain, c_int), cnt, com, off, e))
        im = 0; ia = 0
        do r = 1, nb%type_residue
            if (input%is_active(r) == 1) primary%num_residues(r) = cnt(r)
            do m = 1, cnt(r)
                if (input%is_active(r) == 1) primary%mol_com(:, r, m) = com(3 * im + 1:3 * im + 3)
                im = im + 1
                do a = 1, nb%atom_in_residue(r)
                    if (input%is_active(r) == 1) primary%site_offset(:, r, m, a) = off(3 * ia + 1:3 * ia + 3)
                    ia = ia + 1
                end do
            end do
        end do
        energy%non_coulomb = e(1); energy%coulomb = e(2); energy%recip_coulomb = e(3)
        energy%ewald_self = e(4); energy%intra_coulomb = e(5); energy%total = e(6)
    end subroutine gpu_sync_primary

    subroutine gpu_detach()
        if (c_associated(gpu)) call mmc_destroy(gpu)
        gpu = c_null_ptr
    end subroutine gpu_detach

end module maniac_gpu_glue
