"""spatial_queue_b200 -- B200-native hot paths of cb-cities/spatial_queue.

HP-2  batched shortest-path trees behind the `sp` C ABI      -> sp_interface
HP-1  the 1-second spatial-queue tick (link + node model)    -> tick / queue_class

Everything computes in libsqb200.so (hand-written sm_100a CUDA behind a C ABI, see include/*.h).
"""
__version__ = "0.1.0"
