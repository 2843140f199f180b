"""ctypes binding of libseqm_b200.so (include/seqm_b200.h).  There is no fallback: if the library is
missing or a call fails, an exception is raised."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('SEQM_B200_LIB') or os.path.join(_HERE, 'libseqm_b200.so')      # env: instrumented builds

c_i32p = ctypes.POINTER(ctypes.c_int32)
c_f64p = ctypes.POINTER(ctypes.c_double)
c_u8p = ctypes.POINTER(ctypes.c_uint8)
VP = ctypes.c_void_p


class Topology(ctypes.Structure):
    _fields_ = [('nmol', ctypes.c_int64), ('molsize', ctypes.c_int64), ('natoms', ctypes.c_int64),
                ('npairs', ctypes.c_int64),
                ('Z', VP), ('atom_mol', VP), ('atom_slot', VP), ('idxi', VP), ('idxj', VP), ('pair_cls', VP),
                ('pair_slot', VP), ('cls_pairs', VP * 3), ('ncls', ctypes.c_int64 * 3),
                ('start_i', VP), ('start_j', VP), ('perm_j', VP)]


class WStore(ctypes.Structure):
    _fields_ = [('wHH', VP), ('wXH', VP), ('wXX', VP)]


class ScfOptions(ctypes.Structure):
    _fields_ = [('converger', ctypes.c_int), ('alpha', ctypes.c_double), ('use_sp2', ctypes.c_int),
                ('sp2_eps', ctypes.c_double), ('eps', ctypes.c_double), ('max_iter', ctypes.c_int),
                ('debug', ctypes.c_int), ('async_status', ctypes.c_int)]


class SeqmError(RuntimeError):
    pass


_lib = None


def lib():
    """The loaded library; raises if it has not been built (python __graft_entry__.py / make -C csrc)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SeqmError("libseqm_b200.so not found at %s: build it with `make -C pyseqm_b200/csrc` "
                            "(there is no CPU fallback)" % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        L.seqm_last_error.restype = ctypes.c_char_p
        L.seqm_version.restype = ctypes.c_int
        L.seqm_launch_count.restype = ctypes.c_ulonglong
        L.seqm_sp2_step_count.restype = ctypes.c_ulonglong
        L.seqm_profile_enable.argtypes = [ctypes.c_int]
        L.seqm_profile_enable.restype = None
        L.seqm_profile_report.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
        L.seqm_atompar_bytes.restype = ctypes.c_size_t
        L.seqm_eigh_max_m.restype = ctypes.c_int
        L.seqm_eigh_force_block.argtypes = [ctypes.c_int]
        L.seqm_eigh_force_block.restype = None
        L.seqm_eigh_tolerance.argtypes = [ctypes.c_double]
        L.seqm_eigh_warp_variant.argtypes = [ctypes.c_int]
        L.seqm_eigh_warp_variant.restype = None
        L.seqm_eigh_check_degeneracy.argtypes = [ctypes.c_int]
        L.seqm_eigh_check_degeneracy.restype = None
        L.seqm_pair_grad_reverse.argtypes = [ctypes.c_int]
        L.seqm_pair_grad_reverse.restype = None
        L.seqm_eigh_block_config.argtypes = [ctypes.c_int]
        L.seqm_eigh_block_config.restype = None
        L.seqm_scf_fused.argtypes = [ctypes.c_int]
        L.seqm_scf_fused.restype = None
        L.seqm_eigh_tolerance.restype = None
        L.seqm_sp2_workspace.restype = ctypes.c_size_t
        L.seqm_sp2_workspace.argtypes = [ctypes.c_int64, ctypes.c_int64]
        L.seqm_scf_workspace.restype = ctypes.c_size_t
        L.seqm_scf_workspace.argtypes = [ctypes.c_int64, ctypes.c_int64, ctypes.POINTER(ScfOptions)]
        i64, ci, cd = ctypes.c_int64, ctypes.c_int, ctypes.c_double
        TP = ctypes.POINTER(Topology)
        L.seqm_atom_pack.argtypes = [i64] + [VP] * 13 + [ci, VP, VP]
        L.seqm_hcore.argtypes = [TP, VP, VP, VP, VP, VP, ci, WStore, VP, VP, VP, VP]
        L.seqm_w_expand.argtypes = [TP, WStore, VP, VP]
        L.seqm_fock.argtypes = [TP, WStore] + [VP] * 13
        L.seqm_elec_energy.argtypes = [i64, i64, VP, VP, VP, VP, ci, VP, VP, VP, VP]
        L.seqm_elec_energy_workspace.argtypes = [i64, i64]
        L.seqm_elec_energy_noalloc.argtypes = [i64, i64, VP, VP, VP, VP, ci, VP, VP, VP]
        L.seqm_elec_energy_workspace.restype = ctypes.c_size_t
        L.seqm_eigh_density.argtypes = [i64, i64, ci] + [VP] * 10
        L.seqm_sp2_density.argtypes = [i64, i64, ci, VP, VP, VP, VP, cd, VP, VP, VP, VP, VP]
        L.seqm_sp2_density_async.argtypes = [i64, i64, ci, VP, VP, VP, VP, cd, VP, VP, VP, VP, ci, VP, VP]
        L.seqm_nccl_unique_id.argtypes = [VP]
        L.seqm_nccl_comm_create.argtypes = [VP, ci, ci, ctypes.POINTER(VP)]
        L.seqm_nccl_comm_destroy.argtypes = [VP]
        L.seqm_sp2_nccl_p2p.argtypes = [ci]
        L.seqm_sp2_nccl_p2p.restype = None
        L.seqm_sp2_nccl_subchunks.argtypes = [ci]
        L.seqm_sp2_nccl_subchunks.restype = None
        L.seqm_sp2_shard_plan.argtypes = [ci, ci, ci, ci, c_i32p]
        L.seqm_sp2_shard_plan.restype = None
        L.seqm_sp2_nccl_workspace.restype = ctypes.c_size_t
        L.seqm_sp2_nccl_workspace.argtypes = [i64, i64, ci]
        L.seqm_sp2_density_nccl.argtypes = [VP, ci, i64, i64, ci, VP, VP, VP, VP, cd, VP, VP, VP, VP, VP]
        L.seqm_scf_sp2_comm.argtypes = [VP, ci]
        L.seqm_scf_sp2_comm.restype = None
        L.seqm_scf.argtypes = [TP, WStore] + [VP] * 9 + [ci, ctypes.POINTER(ScfOptions)] + [VP] * 7 + \
                              [VP, VP]
        L.seqm_pair_grad.argtypes = [TP, VP, VP, VP, ci] + [VP] * 7
        L.seqm_dgemm_batched.argtypes = [i64, i64, VP, VP, VP, i64, i64, VP, VP]
        L.seqm_pair_geometry.argtypes = [TP, VP, cd, VP, VP, VP]
        L.seqm_pair_geometry_grad.argtypes = [TP, VP, VP, VP, VP, cd, VP, VP]
        L.seqm_parser_count.argtypes = [i64, i64, VP, VP, cd, cd, VP, VP]
        L.seqm_parser_fill.argtypes = [i64, i64, VP, VP, cd, cd] + [VP] * 24
        L.seqm_xl_extrapolate.argtypes = [i64, ci, cd, c_f64p, VP, i64, VP, VP, VP, VP]
        L.seqm_verlet_kick_drift.argtypes = [i64, cd, VP, VP, VP, VP]
        L.seqm_verlet_finish.argtypes = [i64, cd, cd, VP, VP, VP, VP, VP, VP, VP]
        L.seqm_energy_finish.argtypes = [i64] + [VP] * 8
        L.seqm_md_thermo.argtypes = [i64, i64, cd, cd] + [VP] * 11
        for name in EXPORTS:
            fn = getattr(L, name)
            if name not in ('seqm_last_error', 'seqm_atompar_bytes', 'seqm_sp2_workspace', 'seqm_scf_workspace',
                            'seqm_launch_count', 'seqm_sp2_step_count', 'seqm_eigh_check_degeneracy', 'seqm_profile_enable', 'seqm_eigh_force_block', 'seqm_eigh_tolerance', 'seqm_scf_fused',
                            'seqm_sp2_nccl_subchunks', 'seqm_sp2_shard_plan', 'seqm_elec_energy_workspace', 'seqm_sp2_nccl_p2p', 'seqm_eigh_block_config', 'seqm_pair_grad_reverse', 'seqm_sp2_nccl_workspace', 'seqm_scf_sp2_comm', 'seqm_eigh_warp_variant'):
                fn.restype = ctypes.c_int
        if os.environ.get('SEQM_EIGH_CFG'):
            L.seqm_eigh_block_config(int(os.environ['SEQM_EIGH_CFG']))
        if os.environ.get('SEQM_EIGH_WARP_VARIANT'):
            L.seqm_eigh_warp_variant(int(os.environ['SEQM_EIGH_WARP_VARIANT']))
        if os.environ.get('SEQM_EIGH_TOL'):
            L.seqm_eigh_tolerance(float(os.environ['SEQM_EIGH_TOL']))
        _lib = L
    return _lib


EXPORTS = ['seqm_last_error', 'seqm_version', 'seqm_launch_count', 'seqm_profile_enable', 'seqm_profile_report', 'seqm_atompar_bytes', 'seqm_atom_pack', 'seqm_hcore', 'seqm_w_expand',
           'seqm_fock', 'seqm_elec_energy', 'seqm_eigh_max_m', 'seqm_eigh_force_block', 'seqm_eigh_tolerance', 'seqm_eigh_density', 'seqm_sp2_workspace',
           'seqm_sp2_density', 'seqm_sp2_density_async', 'seqm_sp2_step_count', 'seqm_scf_workspace', 'seqm_scf_fused', 'seqm_scf', 'seqm_pair_grad', 'seqm_dgemm_batched', 'seqm_pair_geometry', 'seqm_pair_geometry_grad',
           'seqm_nccl_unique_id', 'seqm_nccl_comm_create', 'seqm_nccl_comm_destroy', 'seqm_sp2_nccl_subchunks', 'seqm_sp2_shard_plan',
           'seqm_sp2_nccl_workspace', 'seqm_sp2_density_nccl', 'seqm_scf_sp2_comm', 'seqm_elec_energy_workspace', 'seqm_sp2_nccl_p2p', 'seqm_eigh_block_config', 'seqm_pair_grad_reverse',
           'seqm_elec_energy_noalloc', 'seqm_eigh_check_degeneracy', 'seqm_eigh_check_degeneracy_max_m', 'seqm_parser_count', 'seqm_parser_fill', 'seqm_xl_extrapolate', 'seqm_verlet_kick_drift', 'seqm_verlet_finish', 'seqm_energy_finish', 'seqm_md_thermo', 'seqm_eigh_warp_variant']


def check(rc, what):
    if rc != 0:
        msg = lib().seqm_last_error()
        raise SeqmError("%s failed (%d): %s" % (what, rc, msg.decode() if msg else ''))
