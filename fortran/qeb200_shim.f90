!
! qeb200_shim.f90 - ISO_C_BINDING shim between PWscf and libqeb200 (include/qeb200.h).
!
! Drop-in use: compile this file into PW/src (see INTEGRATION.md) and pass
! h_psi_qeb to the solvers where c_bands.f90 passes h_psi / h_psi_gpu
! (PW/src/c_bands.f90:253-254,557,573,850,866), or rename it to h_psi to
! replace the symbol for every caller. The argument list is the reference's
! (PW/src/h_psi.f90:9): all arguments by reference, psi/hpsi contiguous
! COMPLEX(DP)(lda*npol, m).
!
! Not compiled in this repository's CI: neither the build image nor the GPU
! box has a Fortran compiler. The same call sequence is exercised through the
! Python mirror q-e_b200/hpsi.py by tests/test_gpu_parity.py.
!
MODULE qeb200_c
  USE, INTRINSIC :: ISO_C_BINDING
  IMPLICIT NONE
  INTEGER(C_INT), PARAMETER :: QEB_FFT_RHO = 0, QEB_FFT_WAVE = 1
  INTERFACE
    INTEGER(C_INT) FUNCTION qeb_create(ctx, device, stream) BIND(C, name='qeb_create')
      IMPORT; TYPE(C_PTR) :: ctx; INTEGER(C_INT), VALUE :: device; TYPE(C_PTR), VALUE :: stream
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_destroy(ctx) BIND(C, name='qeb_destroy')
      IMPORT; TYPE(C_PTR), VALUE :: ctx
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_pin_host(ctx, p, bytes) BIND(C, name='qeb_pin_host')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, p; INTEGER(C_SIZE_T), VALUE :: bytes
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_unpin_host(ctx, p) BIND(C, name='qeb_unpin_host')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, p
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_device_count() BIND(C, name='qeb_device_count')
      IMPORT
    END FUNCTION
    TYPE(C_PTR) FUNCTION qeb_last_error(ctx) BIND(C, name='qeb_last_error')
      IMPORT; TYPE(C_PTR), VALUE :: ctx
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_set_grid(ctx, nr1, nr2, nr3, nr1x, nr2x, nr3x, ngm, nl, nlm, gamma_only) &
        BIND(C, name='qeb_set_grid')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, nl, nlm
      INTEGER(C_INT), VALUE :: nr1, nr2, nr3, nr1x, nr2x, nr3x, ngm, gamma_only
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_set_kpoint(ctx, npw, npwx, igk, g2kin, vkb, nkb, gstart) BIND(C, name='qeb_set_kpoint')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, igk, g2kin, vkb
      INTEGER(C_INT), VALUE :: npw, npwx, nkb, gstart
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_set_vloc(ctx, vrs) BIND(C, name='qeb_set_vloc')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, vrs
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_set_nonlocal(ctx, nat, nhm, nh_of_atom, ofsbeta, deeq) BIND(C, name='qeb_set_nonlocal')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, nh_of_atom, ofsbeta, deeq
      INTEGER(C_INT), VALUE :: nat, nhm
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_h_psi(ctx, lda, n, m, psi, hpsi) BIND(C, name='qeb_h_psi')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, psi, hpsi; INTEGER(C_INT), VALUE :: lda, n, m
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_h_psi_bgrp(ctx, lda, n, m, psi, hpsi, nbgrp, my_bgrp) BIND(C, name='qeb_h_psi_bgrp')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, psi, hpsi; INTEGER(C_INT), VALUE :: lda, n, m, nbgrp, my_bgrp
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_vloc_psi_k(ctx, lda, n, m, psi, v, hpsi) BIND(C, name='qeb_vloc_psi_k')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, psi, v, hpsi; INTEGER(C_INT), VALUE :: lda, n, m
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_vloc_psi_nc(ctx, lda, n, m, psi, v4, hpsi) BIND(C, name='qeb_vloc_psi_nc')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, psi, v4, hpsi; INTEGER(C_INT), VALUE :: lda, n, m
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_vloc_psi_gamma(ctx, lda, n, m, psi, v, hpsi) BIND(C, name='qeb_vloc_psi_gamma')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, psi, v, hpsi; INTEGER(C_INT), VALUE :: lda, n, m
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_fft(ctx, sign, kind, f, howmany) BIND(C, name='qeb_fft')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, f; INTEGER(C_INT), VALUE :: sign, kind, howmany
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_calbec(ctx, n, beta, nkb, psi, lda, m, becp) BIND(C, name='qeb_calbec')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, beta, psi, becp; INTEGER(C_INT), VALUE :: n, nkb, lda, m
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_add_vuspsi(ctx, lda, n, m, hpsi, becp) BIND(C, name='qeb_add_vuspsi')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, hpsi, becp; INTEGER(C_INT), VALUE :: lda, n, m
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_get_becp(ctx, becp, m) BIND(C, name='qeb_get_becp')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, becp; INTEGER(C_INT), VALUE :: m
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_set_vloc_nc(ctx, v4) BIND(C, name='qeb_set_vloc_nc')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, v4
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_set_nonlocal_nc(ctx, deeq_nc) BIND(C, name='qeb_set_nonlocal_nc')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, deeq_nc
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_h_psi_nc(ctx, lda, n, m, psi, hpsi) BIND(C, name='qeb_h_psi_nc')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, psi, hpsi; INTEGER(C_INT), VALUE :: lda, n, m
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_set_qq_nc(ctx, qq_nc) BIND(C, name='qeb_set_qq_nc')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, qq_nc
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_s_psi_nc(ctx, lda, n, m, psi, spsi) BIND(C, name='qeb_s_psi_nc')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, psi, spsi; INTEGER(C_INT), VALUE :: lda, n, m
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_set_qq(ctx, nat, nhm, qq_at) BIND(C, name='qeb_set_qq')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, qq_at; INTEGER(C_INT), VALUE :: nat, nhm
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_set_grid_dist(ctx, nr1, nr2, nr3, nproc, mype, nr3p, nsw, iss, ismap, nismap, ngm, nl, nlm, &
                                              gamma_only) BIND(C, name='qeb_set_grid_dist')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, nr3p, nsw, iss, ismap, nl, nlm
      INTEGER(C_INT), VALUE :: nr1, nr2, nr3, nproc, mype, nismap, ngm, gamma_only
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_nccl_unique_id(id128) BIND(C, name='qeb_nccl_unique_id')
      IMPORT; TYPE(C_PTR), VALUE :: id128
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_nccl_init(ctx, role, id128, nranks, rank) BIND(C, name='qeb_nccl_init')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, id128; INTEGER(C_INT), VALUE :: role, nranks, rank
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_p2p_export(ctx, blob) BIND(C, name='qeb_p2p_export')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, blob
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_p2p_import(ctx, blobs, nranks) BIND(C, name='qeb_p2p_import')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, blobs; INTEGER(C_INT), VALUE :: nranks
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_hs_psi(ctx, lda, n, m, psi, hpsi, spsi) BIND(C, name='qeb_hs_psi')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, psi, hpsi, spsi; INTEGER(C_INT), VALUE :: lda, n, m
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_cegterg(ctx, npw, npwx, nvec, nvecx, npol, evc, ethr, uspp, e, btype, notcnv, lrot, &
                                        dav_iter, nhpsi) BIND(C, name='qeb_cegterg')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, evc, e, btype, notcnv, dav_iter, nhpsi
      INTEGER(C_INT), VALUE :: npw, npwx, nvec, nvecx, npol, uspp, lrot; REAL(C_DOUBLE), VALUE :: ethr
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_regterg(ctx, npw, npwx, nvec, nvecx, evc, ethr, uspp, e, btype, notcnv, lrot, &
                                        dav_iter, nhpsi) BIND(C, name='qeb_regterg')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, evc, e, btype, notcnv, dav_iter, nhpsi
      INTEGER(C_INT), VALUE :: npw, npwx, nvec, nvecx, uspp, lrot; REAL(C_DOUBLE), VALUE :: ethr
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_s_psi(ctx, lda, n, m, psi, spsi) BIND(C, name='qeb_s_psi')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, psi, spsi; INTEGER(C_INT), VALUE :: lda, n, m
    END FUNCTION
    ! --- device-side producers of the per-k-point state (SURVEY 8 f2) ---
    INTEGER(C_INT) FUNCTION qeb_set_gvectors(ctx, ngm, g, mill) BIND(C, name='qeb_set_gvectors')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, g, mill; INTEGER(C_INT), VALUE :: ngm
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_g2_kin(ctx, xk, tpiba2, qcutz, ecfixed, q2sigma, g2kin_out) BIND(C, name='qeb_g2_kin')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, xk, g2kin_out; REAL(C_DOUBLE), VALUE :: tpiba2, qcutz, ecfixed, q2sigma
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_set_beta_tables(ctx, nsp, nhm, nbetam, lmaxkb, nh, nhtol, nhtolm, indv, nqx, dq, tab_beta) &
        BIND(C, name='qeb_set_beta_tables')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, nh, nhtol, nhtolm, indv, tab_beta
      INTEGER(C_INT), VALUE :: nsp, nhm, nbetam, lmaxkb, nqx; REAL(C_DOUBLE), VALUE :: dq
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_set_structure(ctx, nat, ityp, tau, nr1, nr2, nr3, eigts1, eigts2, eigts3) &
        BIND(C, name='qeb_set_structure')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, ityp, tau, eigts1, eigts2, eigts3; INTEGER(C_INT), VALUE :: nat, nr1, nr2, nr3
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_init_us_2(ctx, q, tpiba, vkb_out) BIND(C, name='qeb_init_us_2')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, q, vkb_out; REAL(C_DOUBLE), VALUE :: tpiba
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_sum_band_rho(ctx, lda, n, m, evc, w, rho) BIND(C, name='qeb_sum_band_rho')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, evc, w, rho; INTEGER(C_INT), VALUE :: lda, n, m
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_sum_band_rho_nc(ctx, lda, n, m, evc, w, rho, domag) BIND(C, name='qeb_sum_band_rho_nc')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, evc, w, rho; INTEGER(C_INT), VALUE :: lda, n, m, domag
    END FUNCTION
    ! --- block-Davidson building blocks (SURVEY 8 f3) ---
    INTEGER(C_INT) FUNCTION qeb_set_g_psi_diag(ctx, n, h_diag, s_diag) BIND(C, name='qeb_set_g_psi_diag')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, h_diag, s_diag; INTEGER(C_INT), VALUE :: n
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_g_psi(ctx, lda, n, m, npol, psi, e) BIND(C, name='qeb_g_psi')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, psi, e; INTEGER(C_INT), VALUE :: lda, n, m, npol
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_col_dot(ctx, n, m, a, lda, b, ldb, res) BIND(C, name='qeb_col_dot')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, a, b, res; INTEGER(C_INT), VALUE :: n, m, lda, ldb
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_dav_gemm_cn(ctx, n, a, lda, ka, b, ldb, mb, c, ldc) BIND(C, name='qeb_dav_gemm_cn')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, a, b, c; INTEGER(C_INT), VALUE :: n, lda, ka, ldb, mb, ldc
    END FUNCTION
    INTEGER(C_INT) FUNCTION qeb_dav_gemm_nn(ctx, n, a, lda, k, b, ldb, m, c, ldc, accumulate) BIND(C, name='qeb_dav_gemm_nn')
      IMPORT; TYPE(C_PTR), VALUE :: ctx, a, b, c; INTEGER(C_INT), VALUE :: n, lda, k, ldb, m, ldc, accumulate
    END FUNCTION
  END INTERFACE
END MODULE qeb200_c

!----------------------------------------------------------------------------
MODULE qeb200_state
  !! One context per MPI rank (= per device, Modules/mp_world.f90:94) and the
  !! "dirty" flags that replace the reference's implicit module state.
  USE, INTRINSIC :: ISO_C_BINDING
  USE qeb200_c
  IMPLICIT NONE
  SAVE
  TYPE(C_PTR) :: ctx = C_NULL_PTR
  LOGICAL :: grid_set = .FALSE.
  INTEGER :: k_set = -1          ! current_k the context holds
  INTEGER :: vloc_version = -1   ! bumped by qeb200_touch_vloc (call it after set_vrs, electrons.f90:1038)
  INTEGER :: vloc_loaded = -2
  INTEGER :: deeq_version = -1   ! bumped by qeb200_touch_deeq (call it after newd, electrons.f90:1045)
  INTEGER :: deeq_loaded = -2
  INTEGER :: spin_loaded = -1
  LOGICAL :: qq_loaded = .FALSE.   ! qq_at is filled once per run (upflib/init_us_1.f90)
  INTEGER :: vnc_loaded = -2       ! vloc_version the context's 2x2 potential belongs to
  INTEGER :: dnc_loaded = -2       ! deeq_version the context's deeq_nc belongs to
  LOGICAL :: qnc_loaded = .FALSE.  ! qq_so / qq_at are filled once per run
CONTAINS
  SUBROUTINE qeb200_check(rc, routine)
    INTEGER(C_INT), INTENT(IN) :: rc
    CHARACTER(*), INTENT(IN) :: routine
    CHARACTER(KIND=C_CHAR), POINTER :: s(:)
    CHARACTER(LEN=256) :: msg
    INTEGER :: i
    IF (rc == 0) RETURN
    msg = ' '
    CALL C_F_POINTER(qeb_last_error(ctx), s, [256])
    DO i = 1, 256
      IF (s(i) == C_NULL_CHAR) EXIT
      msg(i:i) = s(i)
    END DO
    CALL errore(routine, TRIM(msg), INT(rc))   ! UtilXlib/error_handler.f90:10 -> mp_abort
  END SUBROUTINE
  FUNCTION qeb200_local_device() RESULT(dev)
    !! The device mp_world_start binds this rank to: mod(node-local rank, device count)
    !! (Modules/mp_world.f90:84-94: mp_count_nodes gives `key`, cudaSetDevice(mod(key, ndev))).
    USE mp_world, ONLY: world_comm
    USE mp,       ONLY: mp_count_nodes
    INTEGER(C_INT) :: dev
    INTEGER :: nnode_, color, key, ndev
    CALL mp_count_nodes(nnode_, color, key, world_comm)
    ndev = qeb_device_count()
    IF (ndev < 1) CALL errore('qeb200', 'no CUDA device (the library has no CPU fallback)', 1)
    dev = INT(MOD(key, ndev), C_INT)
  END FUNCTION
  SUBROUTINE qeb200_touch_vloc();  vloc_version = vloc_version + 1; END SUBROUTINE
  SUBROUTINE qeb200_touch_deeq();  deeq_version = deeq_version + 1; END SUBROUTINE
  SUBROUTINE qeb200_touch_kpoint(); k_set = -1; END SUBROUTINE   ! after init_us_2 when atoms move

  SUBROUTINE qeb200_rg_communicator()
    !! NCCL communicator of the R&G group (intra_bgrp_comm = dffts%comm): rank 0 makes the id, MPI broadcasts it
    !! (qeb200.h: qeb_nccl_unique_id + qeb_nccl_init); used by the stick<->plane exchange fallback and by the becp
    !! all-reduce (Modules/becmod.f90:395-399)
    USE fft_base, ONLY: dffts
    USE mp,       ONLY: mp_bcast
    INTEGER(C_INT), TARGET :: id(32)          ! 128 bytes
    id = 0
    IF (dffts%mype == 0) CALL qeb200_check(qeb_nccl_unique_id(C_LOC(id)), 'qeb_nccl_unique_id')
    CALL mp_bcast(id, 0, dffts%comm)
    CALL qeb200_check(qeb_nccl_init(ctx, 1_C_INT, C_LOC(id), dffts%nproc, dffts%mype), 'qeb_nccl_init')   ! 1 = QEB_COMM_RG
  END SUBROUTINE
  SUBROUTINE qeb200_rg_peer_mapping()
    !! Peer-to-peer exchange over NVLink for the ranks of one box (qeb_p2p_export -> all-gather of the 256-byte blobs in
    !! rank order -> qeb_p2p_import -> barrier). QEB_P2P=0 in the environment keeps the NCCL exchange.
    USE fft_base, ONLY: dffts
    USE mp,       ONLY: mp_barrier
#if defined(__MPI)
    USE mpi
#endif
    INTEGER(C_INT8_T), ALLOCATABLE, TARGET :: blob(:), blobs(:)
    INTEGER :: ierr
    IF (dffts%nproc < 2) RETURN
    ALLOCATE(blob(256), blobs(256 * dffts%nproc))
    CALL qeb200_check(qeb_p2p_export(ctx, C_LOC(blob)), 'qeb_p2p_export')
#if defined(__MPI)
    CALL MPI_ALLGATHER(blob, 256, MPI_BYTE, blobs, 256, MPI_BYTE, dffts%comm, ierr)
    IF (ierr /= 0) CALL errore('qeb200', 'MPI_Allgather of the peer blobs failed', ierr)
#else
    blobs(1:256) = blob
#endif
    CALL qeb200_check(qeb_p2p_import(ctx, C_LOC(blobs), dffts%nproc), 'qeb_p2p_import')
    CALL mp_barrier(dffts%comm)
    DEALLOCATE(blob, blobs)
  END SUBROUTINE
  SUBROUTINE qeb200_sync_state()
    !! Pushes whatever module state changed since the last call into the context.
    USE fft_base,      ONLY: dffts
    USE gvect,         ONLY: gstart
    USE gvecs,         ONLY: ngms
    USE klist,         ONLY: ngk, igk_k
    USE wvfct,         ONLY: npwx, g2kin, current_k
    USE scf,           ONLY: vrs
    USE lsda_mod,      ONLY: current_spin
    USE uspp,          ONLY: vkb, nkb, deeq, ofsbeta, qq_at, okvan
    USE uspp_param,    ONLY: nh, nhm
    USE ions_base,     ONLY: nat, ityp
    USE control_flags, ONLY: gamma_only
    INTEGER(C_INT), ALLOCATABLE, TARGET :: nh_at(:), ofs0(:)
    INTEGER :: na, ig
    TYPE(C_PTR) :: pnlm, pigk
    IF (.NOT. C_ASSOCIATED(ctx)) CALL qeb200_check(qeb_create(ctx, qeb200_local_device(), C_NULL_PTR), 'qeb_create')
    IF (.NOT. grid_set) THEN
      pnlm = C_NULL_PTR
      IF (gamma_only) pnlm = C_LOC(dffts%nlm)
      ig = 0; IF (gamma_only) ig = 1
      IF (dffts%lpara) THEN
        ! slab-distributed dffts (fft_types.f90:35-185): the rank's descriptor + the R&G communicator (dffts%comm)
        CALL qeb200_check(qeb_set_grid_dist(ctx, dffts%nr1, dffts%nr2, dffts%nr3, dffts%nproc, dffts%mype, &
                          C_LOC(dffts%nr3p), C_LOC(dffts%nsw), C_LOC(dffts%iss), C_LOC(dffts%ismap), SIZE(dffts%ismap), &
                          ngms, C_LOC(dffts%nl), pnlm, INT(ig, C_INT)), 'qeb_set_grid_dist')
        CALL qeb200_rg_communicator()
      ELSE
        CALL qeb200_check(qeb_set_grid(ctx, dffts%nr1, dffts%nr2, dffts%nr3, dffts%nr1x, dffts%nr2x, dffts%nr3x, &
                          ngms, C_LOC(dffts%nl), pnlm, INT(ig, C_INT)), 'qeb_set_grid')
      END IF
      grid_set = .TRUE.
    END IF
    IF (k_set /= current_k) THEN
      pigk = C_NULL_PTR
      IF (.NOT. gamma_only) pigk = C_LOC(igk_k(1, current_k))
      CALL qeb200_check(qeb_set_kpoint(ctx, ngk(current_k), npwx, pigk, C_LOC(g2kin), C_LOC(vkb), nkb, gstart), &
                        'qeb_set_kpoint')
      k_set = current_k
      ! the exchange buffers are sized by qeb_set_kpoint: (re)publish them to the peers of the R&G group
      IF (dffts%lpara) CALL qeb200_rg_peer_mapping()
    END IF
    IF (vloc_loaded /= vloc_version .OR. spin_loaded /= current_spin) THEN
      CALL qeb200_check(qeb_set_vloc(ctx, C_LOC(vrs(1, current_spin))), 'qeb_set_vloc')
      vloc_loaded = vloc_version
    END IF
    IF (nkb > 0 .AND. (deeq_loaded /= deeq_version .OR. spin_loaded /= current_spin)) THEN
      ALLOCATE(nh_at(nat), ofs0(nat))
      DO na = 1, nat
        nh_at(na) = nh(ityp(na)); ofs0(na) = ofsbeta(na)     ! ofsbeta is already a 0-based offset (upflib/uspp.f90)
      END DO
      CALL qeb200_check(qeb_set_nonlocal(ctx, nat, nhm, C_LOC(nh_at), C_LOC(ofs0), C_LOC(deeq(1, 1, 1, current_spin))), &
                        'qeb_set_nonlocal')
      deeq_loaded = deeq_version
      DEALLOCATE(nh_at, ofs0)
      IF (okvan .AND. .NOT. qq_loaded) THEN   ! S = 1 + sum |beta> qq <beta| (PW/src/s_psi.f90:10)
        CALL qeb200_check(qeb_set_qq(ctx, nat, nhm, C_LOC(qq_at)), 'qeb_set_qq')
        qq_loaded = .TRUE.
      END IF
    END IF
    spin_loaded = current_spin
  END SUBROUTINE
  SUBROUTINE qeb200_sync_state_nc()
    !! Magnetic / spin-orbit spinors: the 2x2 potential vrs(:,1:4) (only vrs(:,1) without domag, vloc_psi_acc.f90:329-348)
    !! and deeq_nc (newd_nc / newd_so, PW/src/newd.f90) on top of qeb200_sync_state.
    USE kinds,            ONLY: DP
    USE fft_base,         ONLY: dffts
    USE scf,              ONLY: vrs
    USE uspp,             ONLY: nkb, deeq_nc
    USE noncollin_module, ONLY: domag
    REAL(DP), ALLOCATABLE, TARGET :: v4(:, :)
    CALL qeb200_sync_state()
    IF (vnc_loaded /= vloc_version) THEN
      ALLOCATE (v4(dffts%nnr, 4))
      v4 = 0.0_DP
      v4(:, 1) = vrs(1:dffts%nnr, 1)
      IF (domag) v4(:, 2:4) = vrs(1:dffts%nnr, 2:4)
      CALL qeb200_check(qeb_set_vloc_nc(ctx, C_LOC(v4)), 'qeb_set_vloc_nc')
      DEALLOCATE (v4)
      vnc_loaded = vloc_version
    END IF
    IF (nkb > 0 .AND. dnc_loaded /= deeq_version) THEN
      CALL qeb200_check(qeb_set_nonlocal_nc(ctx, C_LOC(deeq_nc)), 'qeb_set_nonlocal_nc')
      dnc_loaded = deeq_version
    END IF
  END SUBROUTINE
  SUBROUTINE qeb200_sync_qq_nc()
    !! per-atom 2x2-spin overlap of s_psi_nc (PW/src/s_psi.f90:326-360): qq_so(:,:,1:4,nt) with spin-orbit, else qq_at
    !! on the spin diagonal; only ultrasoft types contribute (:328)
    USE kinds,            ONLY: DP
    USE uspp,             ONLY: qq_at, qq_so, okvan
    USE uspp_param,       ONLY: nhm, upf
    USE ions_base,        ONLY: nat, ityp
    USE noncollin_module, ONLY: lspinorb
    COMPLEX(DP), ALLOCATABLE, TARGET :: qq_nc(:, :, :, :)
    INTEGER :: na, nt
    IF (qnc_loaded .OR. .NOT. okvan) RETURN
    ALLOCATE (qq_nc(nhm, nhm, nat, 4))
    qq_nc = (0.0_DP, 0.0_DP)
    DO na = 1, nat
      nt = ityp(na)
      IF (.NOT. upf(nt)%tvanp) CYCLE
      IF (lspinorb) THEN
        qq_nc(:, :, na, 1:4) = qq_so(:, :, 1:4, nt)
      ELSE
        qq_nc(:, :, na, 1) = qq_at(:, :, na)
        qq_nc(:, :, na, 4) = qq_at(:, :, na)
      END IF
    END DO
    CALL qeb200_check(qeb_set_qq_nc(ctx, C_LOC(qq_nc)), 'qeb_set_qq_nc')
    DEALLOCATE (qq_nc)
    qnc_loaded = .TRUE.
  END SUBROUTINE
  SUBROUTINE qeb200_sync_uspp(structure_changed)
    !! Pushes the run-constant inputs of init_us_2 / g2_kin (G-vector list, beta interpolation tables: once) and the
    !! structure-factor phases (after every struc_fact, PW/src/hinit0.f90 / hinit1.f90).
    USE gvect,      ONLY: ngm, g, mill, eigts1, eigts2, eigts3
    USE fft_base,   ONLY: dfftp
    USE ions_base,  ONLY: nat, ityp, tau, nsp
    USE uspp,       ONLY: nhtol, nhtolm, indv
    USE uspp_param, ONLY: nh, nhm, nbetam, lmaxkb
    USE beta_mod,   ONLY: tab_beta, dq, nqx
    LOGICAL, INTENT(IN) :: structure_changed
    LOGICAL, SAVE :: tables_loaded = .FALSE.
    IF (.NOT. C_ASSOCIATED(ctx)) CALL qeb200_sync_state()
    IF (.NOT. tables_loaded) THEN
      CALL qeb200_check(qeb_set_gvectors(ctx, ngm, C_LOC(g), C_LOC(mill)), 'qeb_set_gvectors')
      CALL qeb200_check(qeb_set_beta_tables(ctx, nsp, nhm, nbetam, lmaxkb, C_LOC(nh), C_LOC(nhtol), C_LOC(nhtolm), &
                        C_LOC(indv), nqx, dq, C_LOC(tab_beta)), 'qeb_set_beta_tables')
      tables_loaded = .TRUE.
    END IF
    IF (structure_changed) THEN
      CALL qeb200_check(qeb_set_structure(ctx, nat, C_LOC(ityp), C_LOC(tau), dfftp%nr1, dfftp%nr2, dfftp%nr3, &
                        C_LOC(eigts1), C_LOC(eigts2), C_LOC(eigts3)), 'qeb_set_structure')
      k_set = -1
    END IF
  END SUBROUTINE
  LOGICAL FUNCTION qeb200_hpsi_supported()
    !! .FALSE. for the h_psi_ branches the library does not build (SURVEY 8a a18): they keep the reference routines
    USE realus,   ONLY: real_space
    USE fft_base, ONLY: dffts
    USE ldaU,     ONLY: lda_plus_u
    USE bp,       ONLY: lelfield
    USE xc_lib,   ONLY: exx_is_active, xclib_dft_is
    ! R&G-distributed grids: the slab decomposition (QE's default) is bound through qeb_set_grid_dist below; pencil
    ! decomposition and task groups keep the reference routines
    qeb200_hpsi_supported = .NOT. (real_space .OR. dffts%has_task_groups .OR. lda_plus_u .OR. lelfield .OR. &
                                   exx_is_active() .OR. xclib_dft_is('meta') .OR. &
                                   (dffts%lpara .AND. (dffts%use_pencil_decomposition .OR. dffts%nr1x /= dffts%nr1 .OR. &
                                                       dffts%nr2x /= dffts%nr2)))
  END FUNCTION
  SUBROUTINE qeb200_push_g_psi_diag(n, hd, sd)
    !! g_psi_mod::h_diag / s_diag of the current k-point -> context (PW/src/c_bands.f90 fills them before the solver)
    USE kinds, ONLY: DP
    INTEGER, INTENT(IN) :: n
    REAL(DP), INTENT(IN), TARGET :: hd(*), sd(*)
    CALL qeb200_check(qeb_set_g_psi_diag(ctx, n, C_LOC(hd), C_LOC(sd)), 'qeb_set_g_psi_diag')
  END SUBROUTINE
END MODULE qeb200_state

!----------------------------------------------------------------------------
SUBROUTINE h_psi_qeb(lda, n, m, psi, hpsi)
  !! Same interface as h_psi (PW/src/h_psi.f90:9). Branches the library does not
  !! implement are routed to the original routine (SURVEY 8a a18).
  USE, INTRINSIC :: ISO_C_BINDING
  USE kinds,            ONLY: DP
  USE qeb200_c
  USE qeb200_state
  USE becmod,           ONLY: becp
  USE uspp,             ONLY: nkb
  USE control_flags,    ONLY: gamma_only
  USE noncollin_module, ONLY: npol, noncolin, domag, lspinorb
  USE realus,           ONLY: real_space
  USE fft_base,         ONLY: dffts
  USE ldaU,             ONLY: lda_plus_u
  USE bp,               ONLY: lelfield
  USE xc_lib,           ONLY: exx_is_active, xclib_dft_is
  USE mp_bands,         ONLY: use_bgrp_in_hpsi, inter_bgrp_comm, nbgrp, my_bgrp_id
  USE mp,               ONLY: mp_allgather, mp_size, mp_type_create_column_section, mp_type_free
  IMPLICIT NONE
  INTEGER, INTENT(IN) :: lda, n, m
  COMPLEX(DP), INTENT(IN), TARGET :: psi(lda*npol, m)
  COMPLEX(DP), INTENT(OUT), TARGET :: hpsi(lda*npol, m)
  INTEGER :: m_start, m_end, column_type, m_becp
  INTEGER, ALLOCATABLE :: recv_counts(:), displs(:)
  !
  IF (.NOT. qeb200_hpsi_supported()) THEN
    CALL h_psi(lda, n, m, psi, hpsi)      ! original Fortran path
    RETURN
  END IF
  CALL start_clock('h_psi')
  IF (noncolin .AND. (domag .OR. lspinorb)) THEN
    ! magnetic / spin-orbit spinors: 2x2 potential and spin-dependent deeq_nc (h_psi.f90:184-190,225-229)
    CALL qeb200_sync_state_nc()
    CALL qeb200_check(qeb_h_psi_nc(ctx, lda, n, m, C_LOC(psi), C_LOC(hpsi)), 'h_psi')
    IF (nkb > 0) CALL qeb200_check(qeb_get_becp(ctx, C_LOC(becp%nc), npol*m), 'h_psi')
    CALL stop_clock('h_psi')
    RETURN
  END IF
  CALL qeb200_sync_state()
  m_becp = m
  IF (noncolin) THEN
    ! Non-magnetic spinors without spin-orbit: both components see the same g2kin, the same scalar V(r)
    ! (vloc_psi_nc with domag = .FALSE., vloc_psi_acc.f90:269-369) and deeq_nc(:,:,na,1) = deeq_nc(:,:,na,4) = deeq,
    ! deeq_nc(:,:,na,2:3) = 0 (newd), so H acts on psi(lda*npol, m) exactly as on 2*m columns of leading dimension lda;
    ! becp%nc(nkb,npol,m) has the memory layout of becp(nkb, 2*m).
    CALL qeb200_check(qeb_h_psi(ctx, lda, n, npol*m, C_LOC(psi), C_LOC(hpsi)), 'h_psi')
    IF (nkb > 0) CALL qeb200_check(qeb_get_becp(ctx, C_LOC(becp%nc), npol*m), 'h_psi')
    CALL stop_clock('h_psi')
    RETURN
  END IF
  IF (use_bgrp_in_hpsi .AND. nbgrp > 1 .AND. m > 1) THEN
    ! band groups, h_psi.f90:53-67: own columns on this rank's GPU, then the reference's own all-gather
    ALLOCATE(recv_counts(mp_size(inter_bgrp_comm)), displs(mp_size(inter_bgrp_comm)))
    CALL divide_all(inter_bgrp_comm, m, m_start, m_end, recv_counts, displs)
    CALL qeb200_check(qeb_h_psi_bgrp(ctx, lda, n, m, C_LOC(psi), C_LOC(hpsi), nbgrp, my_bgrp_id), 'h_psi')
    ! like the reference (h_psi.f90:57-62: h_psi_ on psi(1,m_start) with m_end-m_start+1 columns), becp then holds the
    ! projections of THIS group's column block in its first columns, not those of all m bands
    m_becp = MAX(m_end - m_start + 1, 0)
    CALL mp_type_create_column_section(hpsi(1, 1), 0, lda*npol, lda*npol, column_type)
    CALL mp_allgather(hpsi, column_type, recv_counts, displs, inter_bgrp_comm)
    CALL mp_type_free(column_type)
    DEALLOCATE(recv_counts, displs)
  ELSE
    CALL qeb200_check(qeb_h_psi(ctx, lda, n, m, C_LOC(psi), C_LOC(hpsi)), 'h_psi')
  END IF
  ! side effect callers rely on (s_psi.f90:15,132,157): becmod::becp = <beta|psi> of these m columns
  IF (nkb > 0 .AND. m_becp > 0) THEN
    IF (gamma_only) THEN
      CALL qeb200_check(qeb_get_becp(ctx, C_LOC(becp%r), m_becp), 'h_psi')
    ELSE
      CALL qeb200_check(qeb_get_becp(ctx, C_LOC(becp%k), m_becp), 'h_psi')
    END IF
  END IF
  CALL stop_clock('h_psi')
END SUBROUTINE h_psi_qeb

!----------------------------------------------------------------------------
SUBROUTINE vloc_psi_k_qeb(lda, n, m, psi, v, hpsi)
  !! vloc_psi_k (PW/src/vloc_psi_acc.f90:148): hpsi (INOUT) += V_loc psi
  USE, INTRINSIC :: ISO_C_BINDING
  USE kinds, ONLY: DP
  USE qeb200_c; USE qeb200_state
  USE fft_base, ONLY: dffts
  IMPLICIT NONE
  INTEGER, INTENT(IN) :: lda, n, m
  COMPLEX(DP), INTENT(IN), TARGET :: psi(lda, m)
  COMPLEX(DP), INTENT(INOUT), TARGET :: hpsi(lda, m)
  REAL(DP), INTENT(IN), TARGET :: v(dffts%nnr)
  CALL qeb200_sync_state()
  CALL qeb200_check(qeb_vloc_psi_k(ctx, lda, n, m, C_LOC(psi), C_LOC(v), C_LOC(hpsi)), 'vloc_psi_k')
  vloc_loaded = -2   ! the context now holds v, not vrs(:,current_spin)
END SUBROUTINE

SUBROUTINE vloc_psi_gamma_qeb(lda, n, m, psi, v, hpsi)
  !! vloc_psi_gamma (PW/src/vloc_psi_acc.f90:9)
  USE, INTRINSIC :: ISO_C_BINDING
  USE kinds, ONLY: DP
  USE qeb200_c; USE qeb200_state
  USE fft_base, ONLY: dffts
  IMPLICIT NONE
  INTEGER, INTENT(IN) :: lda, n, m
  COMPLEX(DP), INTENT(IN), TARGET :: psi(lda, m)
  COMPLEX(DP), INTENT(INOUT), TARGET :: hpsi(lda, m)
  REAL(DP), INTENT(IN), TARGET :: v(dffts%nnr)
  CALL qeb200_sync_state()
  CALL qeb200_check(qeb_vloc_psi_gamma(ctx, lda, n, m, C_LOC(psi), C_LOC(v), C_LOC(hpsi)), 'vloc_psi_gamma')
  vloc_loaded = -2
END SUBROUTINE

SUBROUTINE vloc_psi_nc_qeb(lda, n, m, psi, v, hpsi)
  !! vloc_psi_nc (PW/src/vloc_psi_acc.f90:269): noncollinear V_loc. domag: the 2x2 product on the device
  !! (qeb_vloc_psi_nc); otherwise v(:,1) on npol*m columns of leading dimension lda (qeb_vloc_psi_k).
  !! v is declared (dfftp%nnr,4) by the reference but only its first dffts%nnr rows are used (:291,:331-347): the four
  !! columns are packed before they go to the library.
  USE, INTRINSIC :: ISO_C_BINDING
  USE kinds, ONLY: DP
  USE qeb200_c; USE qeb200_state
  USE fft_base, ONLY: dffts, dfftp
  USE noncollin_module, ONLY: npol, domag
  IMPLICIT NONE
  INTEGER, INTENT(IN) :: lda, n, m
  COMPLEX(DP), INTENT(IN), TARGET :: psi(lda*npol, m)
  COMPLEX(DP), INTENT(INOUT), TARGET :: hpsi(lda, npol, m)
  REAL(DP), INTENT(IN), TARGET :: v(dfftp%nnr, 4)
  REAL(DP), ALLOCATABLE, TARGET :: v4(:, :)
  IF (dffts%has_task_groups .OR. dffts%lpara) THEN
    CALL vloc_psi_nc(lda, n, m, psi, v, hpsi)   ! reference routine
    RETURN
  END IF
  CALL qeb200_sync_state()
  IF (domag) THEN
    ALLOCATE (v4(dffts%nnr, 4))
    v4(:, :) = v(1:dffts%nnr, 1:4)
    CALL qeb200_check(qeb_vloc_psi_nc(ctx, lda, n, m, C_LOC(psi), C_LOC(v4), C_LOC(hpsi)), 'vloc_psi_nc')
    DEALLOCATE (v4)
  ELSE
    CALL qeb200_check(qeb_vloc_psi_k(ctx, lda, n, npol*m, C_LOC(psi), C_LOC(v), C_LOC(hpsi)), 'vloc_psi_nc')
    vloc_loaded = -2
  END IF
END SUBROUTINE

!----------------------------------------------------------------------------
MODULE qeb200_becmod
  !! Specific procedures to add to the generic calbec of Modules/becmod.f90:40-53
  !! (k and Gamma array forms; the bec_type form dispatches on gamma_only like
  !! calbec_bec_type, becmod.f90:138-172) and the add_vuspsi replacement.
  USE, INTRINSIC :: ISO_C_BINDING
  USE kinds, ONLY: DP
  USE qeb200_c; USE qeb200_state
  IMPLICIT NONE
CONTAINS
  SUBROUTINE calbec_k_qeb(npw, beta, psi, betapsi, nbnd)
    INTEGER, INTENT(IN) :: npw
    COMPLEX(DP), INTENT(IN), TARGET :: beta(:,:), psi(:,:)
    COMPLEX(DP), INTENT(OUT), TARGET :: betapsi(:,:)
    INTEGER, OPTIONAL, INTENT(IN) :: nbnd
    INTEGER :: m
    m = SIZE(psi, 2); IF (PRESENT(nbnd)) m = nbnd
    CALL qeb200_sync_state()
    CALL qeb200_check(qeb_calbec(ctx, npw, C_LOC(beta), SIZE(beta, 2), C_LOC(psi), SIZE(psi, 1), m, C_LOC(betapsi)), 'calbec')
  END SUBROUTINE
  SUBROUTINE calbec_gamma_qeb(npw, beta, psi, betapsi, nbnd)
    INTEGER, INTENT(IN) :: npw
    COMPLEX(DP), INTENT(IN), TARGET :: beta(:,:), psi(:,:)
    REAL(DP), INTENT(OUT), TARGET :: betapsi(:,:)
    INTEGER, OPTIONAL, INTENT(IN) :: nbnd
    INTEGER :: m
    m = SIZE(psi, 2); IF (PRESENT(nbnd)) m = nbnd
    CALL qeb200_sync_state()
    CALL qeb200_check(qeb_calbec(ctx, npw, C_LOC(beta), SIZE(beta, 2), C_LOC(psi), SIZE(psi, 1), m, C_LOC(betapsi)), 'calbec')
  END SUBROUTINE
  SUBROUTINE add_vuspsi_qeb(lda, n, m, hpsi)
    !! add_vuspsi (PW/src/add_vuspsi.f90:10): hpsi += vkb * (D becp), becp = becmod::becp
    USE becmod, ONLY: becp
    USE control_flags, ONLY: gamma_only
    INTEGER, INTENT(IN) :: lda, n, m
    COMPLEX(DP), INTENT(INOUT), TARGET :: hpsi(lda, m)
    CALL qeb200_sync_state()
    IF (gamma_only) THEN
      CALL qeb200_check(qeb_add_vuspsi(ctx, lda, n, m, C_LOC(hpsi), C_LOC(becp%r)), 'add_vuspsi')
    ELSE
      CALL qeb200_check(qeb_add_vuspsi(ctx, lda, n, m, C_LOC(hpsi), C_LOC(becp%k)), 'add_vuspsi')
    END IF
  END SUBROUTINE
END MODULE qeb200_becmod

!----------------------------------------------------------------------------
SUBROUTINE s_psi_qeb(lda, n, m, psi, spsi)
  !! s_psi(lda,n,m,psi,spsi) (PW/src/s_psi.f90:10-76): same EXTERNAL signature cegterg/regterg receive as
  !! s_psi_ptr (KS_Solvers/Davidson/cegterg.f90:110-118,196,467). Uses the projections of these m bands that the
  !! h_psi_qeb call just before left in the context (s_psi.f90:15), so becmod::becp is not re-uploaded.
  !! Spinors (s_psi_nc, :308-370): magnetic / spin-orbit runs use the becp%nc of qeb_h_psi_nc and the 2x2 overlap;
  !! non-magnetic spinors are 2*m scalar columns, as in h_psi_qeb. real_space goes to the reference routine.
  USE, INTRINSIC :: ISO_C_BINDING
  USE kinds,            ONLY: DP
  USE noncollin_module, ONLY: noncolin, npol, domag, lspinorb
  USE realus,           ONLY: real_space
  USE qeb200_c
  USE qeb200_state
  IMPLICIT NONE
  INTEGER, INTENT(IN) :: lda, n, m
  COMPLEX(DP), INTENT(IN),  TARGET :: psi(lda*npol, m)
  COMPLEX(DP), INTENT(OUT), TARGET :: spsi(lda*npol, m)
  IF (real_space) THEN
    CALL s_psi(lda, n, m, psi, spsi)
    RETURN
  END IF
  IF (noncolin .AND. (domag .OR. lspinorb)) THEN
    CALL qeb200_sync_state_nc()
    CALL qeb200_sync_qq_nc()
    CALL qeb200_check(qeb_s_psi_nc(ctx, lda, n, m, C_LOC(psi), C_LOC(spsi)), 's_psi')
    RETURN
  END IF
  CALL qeb200_sync_state()
  CALL qeb200_check(qeb_s_psi(ctx, lda, n, npol*m, C_LOC(psi), C_LOC(spsi)), 's_psi')
END SUBROUTINE s_psi_qeb

!----------------------------------------------------------------------------
SUBROUTINE g_psi_qeb(lda, n, m, npol, psi, e)
  !! g_psi(lda,n,m,npol,psi,e) (PW/src/g_psi.f90:10): the g_psi_ptr of cegterg/regterg (cegterg.f90:110-118,407).
  !! h_diag/s_diag (g_psi_mod) are pushed once per k-point by the caller with qeb_set_g_psi_diag (c_bands.f90 fills
  !! them right before the solver call). psi must be device data (!$acc host_data use_device(psi) in the caller).
  USE, INTRINSIC :: ISO_C_BINDING
  USE kinds, ONLY: DP
  USE qeb200_c
  USE qeb200_state
  IMPLICIT NONE
  INTEGER, INTENT(IN) :: lda, n, m, npol
  COMPLEX(DP), INTENT(INOUT), TARGET :: psi(lda, npol, m)
  REAL(DP), INTENT(IN), TARGET :: e(m)
  IF (npol /= 1) THEN
    ! reached only from the reference solver (the fallback of cegterg_qeb), which has not pushed h_diag(lda, npol);
    ! the library's own npol = 2 solver applies qeb_g_psi itself
    CALL g_psi(lda, n, m, npol, psi, e)
    RETURN
  END IF
  CALL qeb200_check(qeb_g_psi(ctx, lda, n, m, npol, C_LOC(psi), C_LOC(e)), 'g_psi')
END SUBROUTINE g_psi_qeb

!----------------------------------------------------------------------------
SUBROUTINE hs_psi_qeb(lda, n, m, psi, hpsi, spsi)
  !! hs_psi(lda,n,m,psi,hpsi,spsi) (PW/src/hs_psi.f90:7): H and S on the same block, no band-group split; the scalar
  !! (npol = 1) supported branches go to the library in one call, everything else through h_psi_qeb + s_psi_qeb
  USE, INTRINSIC :: ISO_C_BINDING
  USE kinds,            ONLY: DP
  USE noncollin_module, ONLY: npol, noncolin
  USE becmod,           ONLY: becp
  USE uspp,             ONLY: nkb
  USE control_flags,    ONLY: gamma_only
  USE qeb200_c
  USE qeb200_state
  IMPLICIT NONE
  INTEGER, INTENT(IN) :: lda, n, m
  COMPLEX(DP), INTENT(IN),  TARGET :: psi(lda*npol, m)
  COMPLEX(DP), INTENT(OUT), TARGET :: hpsi(lda*npol, m), spsi(lda*npol, m)
  IF (noncolin .OR. .NOT. qeb200_hpsi_supported()) THEN
    CALL h_psi_qeb(lda, n, m, psi, hpsi)
    CALL s_psi_qeb(lda, n, m, psi, spsi)
    RETURN
  END IF
  CALL qeb200_sync_state()
  CALL qeb200_check(qeb_hs_psi(ctx, lda, n, m, C_LOC(psi), C_LOC(hpsi), C_LOC(spsi)), 'hs_psi')
  IF (nkb > 0) THEN   ! becmod::becp as h_psi leaves it
    IF (gamma_only) THEN
      CALL qeb200_check(qeb_get_becp(ctx, C_LOC(becp%r), m), 'hs_psi')
    ELSE
      CALL qeb200_check(qeb_get_becp(ctx, C_LOC(becp%k), m), 'hs_psi')
    END IF
  END IF
END SUBROUTINE hs_psi_qeb

SUBROUTINE hs_1psi_qeb(lda, n, psi, hpsi, spsi)
  !! hs_1psi(lda,n,psi,hpsi,spsi) (PW/src/hs_1psi.f90:7): one band
  USE kinds,            ONLY: DP
  USE noncollin_module, ONLY: npol
  IMPLICIT NONE
  INTEGER, INTENT(IN) :: lda, n
  COMPLEX(DP) :: psi(lda*npol, 1), hpsi(lda*npol, 1), spsi(lda*npol, 1)
  CALL hs_psi_qeb(lda, n, 1, psi, hpsi, spsi)
END SUBROUTINE hs_1psi_qeb

!----------------------------------------------------------------------------
SUBROUTINE cegterg_qeb(h_psi_ptr, s_psi_ptr, uspp, g_psi_ptr, npw, npwx, nvec, nvecx, npol, evc, ethr, e, btype, &
                       notcnv, lrot, dav_iter, nhpsi)
  !! cegterg(...) (KS_Solvers/Davidson/cegterg.f90:18-21), same argument list, for the call in PW/src/c_bands.f90
  !! (diag_bands_k). The whole solve runs on the device with psi, hpsi, spsi resident in HBM (qeb_cegterg); the
  !! EXTERNAL operator arguments are accepted for signature compatibility and used only by the fallback: the library
  !! applies ITS h_psi / s_psi / g_psi, which are what h_psi_qeb / s_psi_qeb / g_psi_qeb would compute. Non-magnetic spinors, more
  !! than one band group, parallel diagonalisation and every h_psi_ branch the library does not build (see h_psi_qeb)
  !! go to the reference cegterg with the pointers the caller passed. g_psi_mod::h_diag / s_diag are pushed here.
  USE, INTRINSIC :: ISO_C_BINDING
  USE kinds,            ONLY: DP
  USE noncollin_module, ONLY: noncolin, domag, lspinorb
  USE mp_bands,         ONLY: nbgrp
  USE g_psi_mod,        ONLY: h_diag, s_diag
  USE qeb200_c
  USE qeb200_state
  IMPLICIT NONE
  EXTERNAL :: h_psi_ptr, s_psi_ptr, g_psi_ptr
  LOGICAL, INTENT(IN) :: uspp, lrot
  LOGICAL :: magnetic
  INTEGER, INTENT(IN) :: npw, npwx, nvec, nvecx, npol
  COMPLEX(DP), INTENT(INOUT), TARGET :: evc(npwx*npol, nvec)
  REAL(DP), INTENT(IN) :: ethr
  REAL(DP), INTENT(OUT), TARGET :: e(nvec)
  INTEGER, INTENT(IN), TARGET :: btype(nvec)
  INTEGER, INTENT(OUT), TARGET :: notcnv, dav_iter, nhpsi
  ! magnetic / spin-orbit spinors: the library's solver runs with npol = 2 on its own h_psi_nc / s_psi_nc / g_psi
  ! (kdim = npwx*npol, cegterg.f90:131-142); non-magnetic spinors keep the reference solver
  magnetic = noncolin .AND. (domag .OR. lspinorb) .AND. npol == 2
  IF ((noncolin .AND. .NOT. magnetic) .OR. (npol /= 1 .AND. .NOT. magnetic) .OR. nbgrp > 1 .OR. &
      .NOT. qeb200_hpsi_supported()) THEN
    CALL cegterg(h_psi_ptr, s_psi_ptr, uspp, g_psi_ptr, npw, npwx, nvec, nvecx, npol, evc, ethr, e, btype, &
                 notcnv, lrot, dav_iter, nhpsi)
    RETURN
  END IF
  IF (magnetic) THEN
    CALL qeb200_sync_state_nc()
    IF (uspp) CALL qeb200_sync_qq_nc()
    CALL qeb200_push_g_psi_diag(npwx*npol, h_diag, s_diag)   ! h_diag(npwx, npol), s_diag(npwx, npol)
  ELSE
    CALL qeb200_sync_state()
    CALL qeb200_push_g_psi_diag(npw, h_diag, s_diag)
  END IF
  CALL qeb200_check(qeb_cegterg(ctx, npw, npwx, nvec, nvecx, npol, C_LOC(evc), ethr, MERGE(1, 0, uspp), C_LOC(e), &
                                C_LOC(btype), C_LOC(notcnv), MERGE(1, 0, lrot), C_LOC(dav_iter), C_LOC(nhpsi)), 'cegterg')
END SUBROUTINE cegterg_qeb

!----------------------------------------------------------------------------
SUBROUTINE regterg_qeb(h_psi_ptr, s_psi_ptr, uspp, g_psi_ptr, npw, npwx, nvec, nvecx, evc, ethr, e, btype, &
                       notcnv, lrot, dav_iter, nhpsi)
  !! regterg(...) (KS_Solvers/Davidson/regterg.f90:18-21), the gamma_only twin of cegterg_qeb (diag_bands_gamma).
  USE, INTRINSIC :: ISO_C_BINDING
  USE kinds,     ONLY: DP
  USE mp_bands,  ONLY: nbgrp
  USE g_psi_mod, ONLY: h_diag, s_diag
  USE qeb200_c
  USE qeb200_state
  IMPLICIT NONE
  EXTERNAL :: h_psi_ptr, s_psi_ptr, g_psi_ptr
  LOGICAL, INTENT(IN) :: uspp, lrot
  INTEGER, INTENT(IN) :: npw, npwx, nvec, nvecx
  COMPLEX(DP), INTENT(INOUT), TARGET :: evc(npwx, nvec)
  REAL(DP), INTENT(IN) :: ethr
  REAL(DP), INTENT(OUT), TARGET :: e(nvec)
  INTEGER, INTENT(IN), TARGET :: btype(nvec)
  INTEGER, INTENT(OUT), TARGET :: notcnv, dav_iter, nhpsi
  IF (nbgrp > 1 .OR. .NOT. qeb200_hpsi_supported()) THEN
    CALL regterg(h_psi_ptr, s_psi_ptr, uspp, g_psi_ptr, npw, npwx, nvec, nvecx, evc, ethr, e, btype, &
                 notcnv, lrot, dav_iter, nhpsi)
    RETURN
  END IF
  CALL qeb200_sync_state()
  CALL qeb200_push_g_psi_diag(npw, h_diag, s_diag)
  CALL qeb200_check(qeb_regterg(ctx, npw, npwx, nvec, nvecx, C_LOC(evc), ethr, MERGE(1, 0, uspp), C_LOC(e), &
                                C_LOC(btype), C_LOC(notcnv), MERGE(1, 0, lrot), C_LOC(dav_iter), C_LOC(nhpsi)), 'regterg')
END SUBROUTINE regterg_qeb

!----------------------------------------------------------------------------
SUBROUTINE init_us_2_qeb(npw_, igk_, q_, vkb_)
  !! init_us_2(npw,igk,q,vkb) (upflib/uspp_init / init_us_2_acc.f90:10) for the k-point the context holds: vkb is built
  !! on the device and stays there for h_psi; vkb_ (module uspp::vkb) receives a copy for the routines outside the
  !! library (forces, stress, sum_bec). qeb_set_beta_tables / qeb_set_gvectors are pushed once per run and
  !! qeb_set_structure after every struc_fact by qeb200_sync_uspp (module qeb200_state).
  USE, INTRINSIC :: ISO_C_BINDING
  USE kinds,     ONLY: DP
  USE cell_base, ONLY: tpiba
  USE wvfct,     ONLY: npwx
  USE uspp,      ONLY: nkb
  USE qeb200_c
  USE qeb200_state
  IMPLICIT NONE
  INTEGER, INTENT(IN) :: npw_, igk_(npw_)
  REAL(DP), INTENT(IN), TARGET :: q_(3)
  COMPLEX(DP), INTENT(OUT), TARGET :: vkb_(npwx, nkb)
  CALL qeb200_sync_state()
  CALL qeb200_check(qeb_init_us_2(ctx, C_LOC(q_), tpiba, C_LOC(vkb_)), 'init_us_2')
END SUBROUTINE init_us_2_qeb

!----------------------------------------------------------------------------
SUBROUTINE invfft_qeb(fft_kind, f, howmany)
  !! invfft(fft_kind, f, dffts, howmany) (FFTXlib/src/fft_interfaces.f90:19-47) for the serial smooth grid
  USE, INTRINSIC :: ISO_C_BINDING
  USE kinds, ONLY: DP
  USE qeb200_c; USE qeb200_state
  IMPLICIT NONE
  CHARACTER(LEN=*), INTENT(IN) :: fft_kind
  COMPLEX(DP), INTENT(INOUT), TARGET :: f(*)
  INTEGER, INTENT(IN) :: howmany
  INTEGER(C_INT) :: kind
  kind = QEB_FFT_RHO; IF (fft_kind == 'Wave') kind = QEB_FFT_WAVE
  CALL qeb200_sync_state()
  CALL qeb200_check(qeb_fft(ctx, 1_C_INT, kind, C_LOC(f), howmany), 'invfft')
END SUBROUTINE

SUBROUTINE fwfft_qeb(fft_kind, f, howmany)
  !! fwfft(fft_kind, f, dffts, howmany) (FFTXlib/src/fft_interfaces.f90:49-80)
  USE, INTRINSIC :: ISO_C_BINDING
  USE kinds, ONLY: DP
  USE qeb200_c; USE qeb200_state
  IMPLICIT NONE
  CHARACTER(LEN=*), INTENT(IN) :: fft_kind
  COMPLEX(DP), INTENT(INOUT), TARGET :: f(*)
  INTEGER, INTENT(IN) :: howmany
  INTEGER(C_INT) :: kind
  kind = QEB_FFT_RHO; IF (fft_kind == 'Wave') kind = QEB_FFT_WAVE
  CALL qeb200_sync_state()
  CALL qeb200_check(qeb_fft(ctx, -1_C_INT, kind, C_LOC(f), howmany), 'fwfft')
END SUBROUTINE
