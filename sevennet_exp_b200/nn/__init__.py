"""``sevenn.nn``-compatible modules on the B200 kernels (the reference's module seam)."""
from .modules import (AtomGraphSequential, AtomReduce, BesselBasis, EdgeEmbedding, EdgePreprocess,  # noqa: F401
                      EquivariantGate, FCN_e3nn, ForceOutput, ForceStressOutput, ForceStressOutputFromEdge, IrrepsConvolution, IrrepsLinear, ModalWiseRescale, MultiCutoffEdgeEmbedding, OnehotEmbedding,
                      PolynomialCutoff, Rescale, SelfConnectionIntro, SelfConnectionLinearIntro,
                      SelfConnectionOutro, SpeciesWiseRescale, SphericalEncoding, XPLORCutoff,
                      build_E3_equivariant_model, init_edge_embedding)
from ._backend import set_lib  # noqa: F401
