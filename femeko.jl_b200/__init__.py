"""femeko.jl_b200 -- B200-native hot path of Femeko.jl (P1 tet assembly, Jacobi-PCG, field recovery).

Host-side mirror of the reference interface (src/FEM3D.jl, src/mesh.jl, src/magneticProperties.jl)
above the C-ABI library ``csrc/libfemeko_cuda.so``.  Imported as ``femeko_jl_b200``.
"""
from .mesh import MESH, kuhn_cube, delaunay_ball, tri_grid, delaunay_disc, to_p2, kuhn_cube_p2, edgeNumbering, edgeDofs, NEDELEC_EDGES, elementVolume, areaTriangle, normalSurface, mu0  # noqa: F401
from .meshio import save2file, saveMesh, loadMesh, read_msh  # noqa: F401
from .lib import FemekoError, build, load  # noqa: F401
from .fem import (Femeko, FemekoMulti, SparseCSR, stiffnessMatrix, massMatrix, stiffnessMatrix2D, massMatrix2D, quadraticStiffnessMatrix, quadraticMassMatrix, quadraticConvectionMatrix, tangentialStiffnessMatrix,  # noqa: F401
                  localStiffnessMatrix, lagrange, BoundaryIntegral, cg, lineSearch, recoverField, Spline1D,
                  residual_norm_noalloc, NOALLOC_MAX_ROWS, nedelecLocalStiffness, nedelecStiffnessMatrix, nedelecTangentialStiffness)
from .materials import (DATA, gradient, materialPermeability, loadMaterial, synthetic_gd_like,  # noqa: F401
                        synthetic_iron_like)
from .magnetostatics import picard, newton_raphson, NonlinearResult, nedelec3D, NedelecResult  # noqa: F401
