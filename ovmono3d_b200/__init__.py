"""B200-native 2D->3D lift for OVMono3D's ROIHeads3DGDINO (ROIAlign -> cube head -> decode)."""
__version__ = "0.1.0"
