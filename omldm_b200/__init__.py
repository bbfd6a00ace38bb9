"""omldm_b200 - B200-native distributed online learning with the capabilities of ArisKonidaris/OMLDM.

Public entry points:
    omldm_b200.engine.Engine          GPU engine (one worker + hub shard per GPU)
    omldm_b200.runtime.job.HostJob    fp64 message-driven host engine (oracle / CPU baseline)
    omldm_b200.cli                    command line (`python -m omldm_b200.cli --help`)
    omldm_b200.stream.ring            pinned host rings and synthetic stream producers
    omldm_b200.ops.functional         one Python function per sm_100a kernel
"""
__version__ = "0.1.0"
