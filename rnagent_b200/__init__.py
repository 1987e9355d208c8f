"""rnagent_b200 -- B200-native reward hot path of RLIF (andriusbern/RNAgent).

Batched Zuker MFE folding + dot-bracket Hamming distance behind the reference's own call
signatures.  Import is cheap and GPU-free; the first fold binds the CUDA library
(rnagent_b200/librnafold_b200.so) and raises if it, or a B200, is missing.
"""
from .dataset import Dataset
from .dotbracket import DotBracket
from .ensemble import EnsembleResult, ensemble_batch, eval_structure_batch
from .environments import RnaDesignBatch, RolloutStats, create_env, rollout, rollout_device, shard_sequences, shard_targets
from .permute import PermuteResult, permute_batch
from .solution import Solution
from .tester import PolicyModel, Tester
from .utils import hamming_distance, load_dataset, load_fasta, load_sequence, read_fasta_records, write_fasta
from .vienna import (PARAM_FILES, PARAMETERS, config_vienna, cvar, fold, fold_batch, fold_hamming_batch,
                     read_parameter_file, set_vienna_params)

__all__ = ["fold", "fold_batch", "fold_hamming_batch", "read_parameter_file", "set_vienna_params",
           "config_vienna", "cvar", "hamming_distance", "PARAMETERS", "PARAM_FILES", "DotBracket", "RnaDesignBatch",
           "RolloutStats", "create_env", "rollout", "rollout_device", "shard_targets", "shard_sequences", "permute_batch", "PermuteResult", "Solution",
           "ensemble_batch", "eval_structure_batch", "EnsembleResult", "Dataset", "Tester", "PolicyModel", "load_sequence", "load_dataset", "load_fasta", "write_fasta", "read_fasta_records"]
