"""hicembler_b200 -- B200-native (sm_100a) hot path of HiCembler: read-pair binning, Sinkhorn-Knopp
balancing, contacts-to-distance transform and NN-chain linkage, behind the reference's own Python
function names.  See DESIGN.md / INTEGRATION.md."""
__version__ = "0.1.0"
