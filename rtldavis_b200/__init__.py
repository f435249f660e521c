"""rtldavis_b200 -- B200-native demodulation hot path of rtldavis.

``dsp`` / ``crc`` / ``protocol`` mirror the reference modules for this path;
``synth`` makes seeded synthetic captures.  Importing the package does not load
the CUDA library; constructing a demodulator does, and fails loudly without it.
"""
__all__ = ["dsp", "crc", "protocol", "synth"]
