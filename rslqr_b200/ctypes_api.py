"""ctypes mirror of the public C structs and prototypes (include/*.h == the reference's src/*.h layouts).

No library is loaded here: ``bind_host_api(lib, solver_struct)`` declares the prototypes on whatever library it is
given -- librslqr.so (rslqr_b200/lib.py) or, in the tests and the CPU reference arm of bench.py, the compiled reference
(oracle/_ref/libref_rslqr.so), which exports the same names. Keeping this table apart from lib.py lets the reference
arm run without mapping the product library at all.
"""
import ctypes as C

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)


class Matrix(C.Structure):
    _fields_ = [("rows", C.c_int), ("cols", C.c_int), ("data", c_dp)]


class LQRData(C.Structure):
    _fields_ = [("nstates", C.c_int), ("ninputs", C.c_int)] + [(k, c_dp) for k in "QRqrcABd"]


class LQRProblem(C.Structure):
    _fields_ = [("nhorizon", C.c_int), ("x0", c_dp), ("lqrdata", C.POINTER(C.POINTER(LQRData)))]


class NdFactor(C.Structure):
    _fields_ = [("lambda_", Matrix), ("state", Matrix), ("input", Matrix)]


class NdData(C.Structure):
    _fields_ = [("nstates", C.c_int), ("ninputs", C.c_int), ("nsegments", C.c_int), ("depth", C.c_int),
                ("width", C.c_int), ("data", c_dp), ("factors", C.POINTER(NdFactor))]


class CholeskyInfo(C.Structure):
    _fields_ = [("uplo", C.c_char), ("success", C.c_int), ("lib", C.c_char), ("fact", C.c_void_p),
                ("is_freed", C.c_int)]


class NdLqrCholeskyFactors(C.Structure):
    _fields_ = [("depth", C.c_int), ("nhorizon", C.c_int), ("cholinfo", C.POINTER(CholeskyInfo)),
                ("numfacts", C.c_int)]


class UnitRange(C.Structure):
    _fields_ = [("start", C.c_int), ("stop", C.c_int)]


class BinaryNode(C.Structure):
    pass


BinaryNode._fields_ = [("idx", C.c_int), ("level", C.c_int), ("levelidx", C.c_int), ("left_inds", UnitRange),
                       ("right_inds", UnitRange), ("parent", C.POINTER(BinaryNode)),
                       ("left_child", C.POINTER(BinaryNode)), ("right_child", C.POINTER(BinaryNode))]


class OrderedBinaryTree(C.Structure):
    _fields_ = [("root", C.POINTER(BinaryNode)), ("node_list", C.POINTER(BinaryNode)),
                ("num_elements", C.c_int), ("depth", C.c_int)]


class NdLqrProfile(C.Structure):
    _fields_ = [(k, C.c_double) for k in ("t_total_ms", "t_leaves_ms", "t_products_ms", "t_cholesky_ms",
                                           "t_cholsolve_ms", "t_shur_ms")] + [("num_threads", C.c_int)]


_REF_SOLVER_FIELDS = [
    ("nstates", C.c_int), ("ninputs", C.c_int), ("nhorizon", C.c_int), ("depth", C.c_int), ("nvars", C.c_int),
    ("tree", OrderedBinaryTree), ("diagonals", C.POINTER(Matrix)), ("data", C.POINTER(NdData)),
    ("fact", C.POINTER(NdData)), ("soln", C.POINTER(NdData)), ("cholfacts", C.POINTER(NdLqrCholeskyFactors)),
    ("solve_time_ms", C.c_double), ("linalg_time_ms", C.c_double), ("profile", NdLqrProfile),
    ("num_threads", C.c_int)]


class RefSolverStruct(C.Structure):  # layout of the reference's NdLqrSolver (src/solver.h:114-130)
    _fields_ = _REF_SOLVER_FIELDS


class SolverStruct(C.Structure):  # this build: same prefix + appended fields (include/solver.h)
    _fields_ = _REF_SOLVER_FIELDS + [("device", C.c_void_p), ("flat_knots", c_dp), ("flat_x0", c_dp),
                                     ("mirror_fact", C.c_int), ("pinned_fact", C.c_int), ("pinned_soln", C.c_int),
                                     ("pinned_knots", C.c_int), ("initialized", C.c_int), ("fact_dirty", C.c_int),
                                     ("status_host", C.c_void_p)]


def bind_host_api(lib, solver_struct):
    """Declare the prototypes of the public C API on `lib` (used for librslqr.so and, in the tests,
    for the compiled reference, which exports the same names)."""
    def sig(name, restype, *argtypes):
        try:
            fn = getattr(lib, name)
        except AttributeError:
            return
        fn.restype = restype
        fn.argtypes = list(argtypes)
    SP = C.POINTER(solver_struct)
    PP = C.POINTER(LQRProblem)
    DP = C.POINTER(LQRData)
    NP = C.POINTER(NdData)
    sig("ndlqr_NewLQRData", DP, C.c_int, C.c_int)
    sig("ndlqr_FreeLQRData", C.c_int, DP)
    sig("ndlqr_InitializeLQRData", C.c_int, DP, c_dp, c_dp, c_dp, c_dp, C.c_double, c_dp, c_dp, c_dp)
    sig("ndlqr_CopyLQRData", C.c_int, DP, DP)
    sig("ndlqr_NewLQRProblem", PP, C.c_int, C.c_int, C.c_int)
    sig("ndlqr_InitializeLQRProblem", C.c_int, PP, c_dp, C.POINTER(DP))
    sig("ndlqr_FreeLQRProblem", C.c_int, PP)
    sig("ndlqr_ReadLQRProblemJSONFile", PP, C.c_char_p)
    sig("ndlqr_ReadLQRDataJSONFile", DP, C.c_char_p)
    sig("ReadMatrixJSONFile", Matrix, C.c_char_p, C.c_char_p)
    sig("FreeMatrix", C.c_int, C.POINTER(Matrix))
    sig("ndlqr_NewNdLqrSolver", SP, C.c_int, C.c_int, C.c_int)
    sig("ndlqr_FreeNdLqrSolver", C.c_int, SP)
    sig("ndlqr_InitializeWithLQRProblem", C.c_int, PP, SP)
    sig("ndlqr_ResetSolver", None, SP)
    sig("ndlqr_Solve", C.c_int, SP)
    sig("ndlqr_SolveLQRProblem", C.c_int, PP, SP)
    sig("ndlqr_SolveRhsOnly", C.c_int, SP)
    sig("ndlqr_SyncFactMirror", C.c_int, SP)
    sig("ndlqr_GetSolution", Matrix, SP)
    sig("ndlqr_CopySolution", C.c_int, SP, c_dp)
    sig("ndlqr_GetNumVars", C.c_int, SP)
    sig("ndlqr_SetNumThreads", C.c_int, SP, C.c_int)
    sig("ndlqr_GetNumThreads", C.c_int, SP)
    sig("ndlqr_GetProfile", NdLqrProfile, SP)
    sig("ndlqr_SolveLeaf", C.c_int, SP, C.c_int)
    sig("ndlqr_SolveLeaves", C.c_int, SP)
    sig("ndlqr_FactorInnerProduct", C.c_int, NP, NP, C.c_int, C.c_int, C.c_int)
    sig("ndlqr_SolveCholeskyFactor", C.c_int, NP, C.POINTER(CholeskyInfo), C.c_int, C.c_int, C.c_int)
    sig("ndlqr_UpdateShurFactor", C.c_int, NP, NP, C.c_int, C.c_int, C.c_int, C.c_int, C.c_bool)
    sig("ndlqr_ShouldCalcLambda", C.c_bool, C.POINTER(OrderedBinaryTree), C.c_int, C.c_int)
    sig("ndlqr_ComputeShurCompliment", C.c_int, SP, C.c_int, C.c_int, C.c_int)
    sig("ndlqr_SolverFactorCholesky", C.c_int, SP, C.c_int, C.c_int)
    sig("ndlqr_NewNdData", NP, C.c_int, C.c_int, C.c_int, C.c_int)
    sig("ndlqr_FreeNdData", C.c_int, NP)
    sig("ndlqr_GetNdFactor", C.c_int, NP, C.c_int, C.c_int, C.POINTER(C.POINTER(NdFactor)))
    sig("ndlqr_ResetNdData", None, NP)
    sig("ndlqr_BuildTree", OrderedBinaryTree, C.c_int)
    sig("ndlqr_FreeTree", C.c_int, C.POINTER(OrderedBinaryTree))
    sig("ndlqr_GetIndexFromLeaf", C.c_int, C.POINTER(OrderedBinaryTree), C.c_int, C.c_int)
    sig("ndlqr_GetIndexLevel", C.c_int, C.POINTER(OrderedBinaryTree), C.c_int)
    sig("ndlqr_GetIndexAtLevel", C.c_int, C.POINTER(OrderedBinaryTree), C.c_int, C.c_int)
    sig("ndlqr_NewCholeskyFactors", C.POINTER(NdLqrCholeskyFactors), C.c_int, C.c_int)
    sig("ndlqr_FreeCholeskyFactors", C.c_int, C.POINTER(NdLqrCholeskyFactors))
    sig("ndlqr_GetQFactorizon", C.c_int, C.POINTER(NdLqrCholeskyFactors), C.c_int, C.POINTER(C.POINTER(CholeskyInfo)))
    sig("ndlqr_GetRFactorizon", C.c_int, C.POINTER(NdLqrCholeskyFactors), C.c_int, C.POINTER(C.POINTER(CholeskyInfo)))
    sig("ndlqr_GetSFactorization", C.c_int, C.POINTER(NdLqrCholeskyFactors), C.c_int, C.c_int,
        C.POINTER(C.POINTER(CholeskyInfo)))
    sig("rslqr_synth_problem", PP, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_int)
    sig("ndlqr_LQRDataBlockSize", C.c_int, C.c_int, C.c_int)
    ip = C.POINTER(C.c_int)
    sig("rslqr_WriteProblemsBinary", C.c_int, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p)
    sig("rslqr_ReadProblemsBinaryHeader", C.c_int, C.c_char_p, ip, ip, ip, ip)
    sig("rslqr_ReadProblemsBinary", C.c_int, C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p)
    sig("rslqr_WriteLQRProblemsBinary", C.c_int, C.c_char_p, C.POINTER(PP), C.c_int)
    sig("rslqr_ReadLQRProblemBinary", PP, C.c_char_p, C.c_int)
    return lib
