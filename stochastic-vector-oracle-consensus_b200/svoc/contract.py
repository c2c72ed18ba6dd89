"""Host-side mirror of the contract ABI `IOracleConsensusNDS` (contract/src/contract.cairo:4-35) over
the single-round engine: same entry points, felt252 encoding and revert behaviour, so the reference's
integration tests (contract/tests/test_contract.cairo) read the same here.

Every `update_prediction` after activation recomputes the whole consensus from the stored vectors,
exactly as the contract does (sequential-commit replay, contract.cairo:331-343, 447-449); the arithmetic
runs in libsvoc (`svoc_consensus_fixed`), bit-exact with the Cairo fixed-point code.  A Cairo panic
becomes a `ContractPanic` and leaves the state untouched (a reverted transaction).  The admin
replacement vote (contract.cairo:547-580, 661-738) is a small state machine kept here on the host.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

from . import engine
from ._lib import STATUS_NAMES
from .client import felt_to_i128, i128_to_felt

WSAD = 1_000_000


class ContractPanic(RuntimeError):
    """A reverted call.  `status` is the SVOC_ST_* code for consensus panics, 0 for plain asserts."""

    def __init__(self, msg: str, status: int = 0):
        super().__init__(msg)
        self.status = status


class OracleConsensusNDS:
    def __init__(self, admins: Sequence, enable_oracle_replacement: bool, required_majority: int,
                 n_failing_oracles: int, constrained: bool, unconstrained_max_spread: int, dimension: int,
                 oracles: Sequence, device=None):
        # constructor, contract.cairo:235-265
        self._device = device                           # resolved at the first consensus (no GPU needed before)
        self.dimension = int(dimension)
        self.admins = list(admins)
        self.oracle_addresses = list(oracles)
        n = len(self.oracle_addresses)
        self.enabled = [False] * n
        self.reliable = [True] * n                      # fill_oracles: reliable = true
        self.values = [[0] * self.dimension for _ in range(n)]
        self.n_active_oracles = 0
        self.enable_oracle_replacement = bool(enable_oracle_replacement)
        self.required_majority = int(required_majority)
        self.n_failing_oracles = int(n_failing_oracles)
        self.constrained = bool(constrained)
        self.unconstrained_max_spread = felt_to_i128(int(unconstrained_max_spread))
        self._stray_votes = {}                          # votes written at receiver keys beyond the admin list
        self._reset_votes()
        self.consensus_value = [0] * self.dimension
        self.reliability_first_pass = 0
        self.reliability_second_pass = 0
        self.skewness = [0] * self.dimension
        self.kurtosis = [0] * self.dimension
        self._active = False
        self.caller = None                              # starknet::testing::set_contract_address

    # ---- test helper mirroring starknet::testing::set_contract_address (tests/test_contract.cairo:108)
    def set_caller(self, address) -> None:
        self.caller = address

    def _reset_votes(self) -> None:
        a = len(self.admins)
        self.vote_matrix = [[False] * a for _ in range(a)]       # [emitter][receiver]
        self.replacement_propositions: List[Optional[Tuple[int, object]]] = [None] * a

    def _admin_index(self) -> int:
        if self.caller not in self.admins:
            raise ContractPanic("not an admin")
        return self.admins.index(self.caller)

    # ---- consensus ---------------------------------------------------------------------------
    def update_prediction(self, prediction: Sequence[int], caller=None) -> None:
        """contract.cairo:588-603.  `prediction` is a FeltVector (felt252-encoded wsad)."""
        caller = self.caller if caller is None else caller
        vec = [felt_to_i128(int(v)) for v in prediction]
        if len(vec) != self.dimension:
            raise ContractPanic("wrong dimension")
        if self.constrained and any(not (0 <= v <= WSAD) for v in vec):
            raise ContractPanic("interval error", 1)          # nd_interval_check, math.cairo:298-310
        if caller not in self.oracle_addresses:
            raise ContractPanic("not an oracle")
        idx = self.oracle_addresses.index(caller)
        values = [list(v) for v in self.values]
        values[idx] = vec
        enabled = list(self.enabled)
        n_active = self.n_active_oracles + (0 if enabled[idx] else 1)
        enabled[idx] = True
        if n_active != len(self.oracle_addresses):             # activation gate, contract.cairo:447-449
            self.values, self.enabled, self.n_active_oracles = values, enabled, n_active
            return
        dev = torch.device("cuda", torch.cuda.current_device()) if self._device is None else torch.device(self._device)
        x = torch.tensor([values], dtype=torch.int64, device=dev)
        r = engine.consensus_fixed(x, self.n_failing_oracles, constrained=self.constrained,
                                   max_spread=self.unconstrained_max_spread, want_moments=True)
        status = int(r.status[0])
        if status != 0:                                         # the transaction reverts: nothing is stored
            raise ContractPanic(f"consensus panicked: {STATUS_NAMES.get(status, status)}", status)
        self.values, self.enabled, self.n_active_oracles = values, enabled, n_active
        self.reliability_first_pass = int(r.rel1[0])
        self.reliability_second_pass = int(r.rel2[0])
        self.reliable = r.reliable[0].tolist()
        self.consensus_value = r.consensus[0].tolist()
        self.skewness = r.skew[0].tolist()
        self.kurtosis = r.kurt[0].tolist()
        self._active = True

    def consensus_active(self) -> bool:
        return self._active

    def get_consensus_value(self) -> List[int]:
        return [i128_to_felt(v) for v in self.consensus_value]

    def get_first_pass_consensus_reliability(self) -> int:
        return i128_to_felt(self.reliability_first_pass)

    def get_second_pass_consensus_reliability(self) -> int:
        return i128_to_felt(self.reliability_second_pass)

    def get_skewness(self) -> List[int]:
        return [i128_to_felt(v) for v in self.skewness]

    def get_kurtosis(self) -> List[int]:
        return [i128_to_felt(v) for v in self.kurtosis]

    def get_predictions_dimension(self) -> int:
        return self.dimension

    def get_admin_list(self) -> list:
        return list(self.admins)

    def get_oracle_list(self) -> list:
        return list(self.oracle_addresses)

    def get_oracle_value_list(self):
        """contract.cairo:772-798: (address, FeltVector, enabled, reliable) per oracle; admins only."""
        if self.caller not in self.admins:
            raise ContractPanic("not admin")
        return [(self.oracle_addresses[i], [i128_to_felt(v) for v in self.values[i]], self.enabled[i], self.reliable[i])
                for i in range(len(self.oracle_addresses))]

    # ---- admin replacement vote (contract.cairo:547-580, 661-738) ------------------------------
    def update_proposition(self, proposition: Optional[Tuple[int, object]]) -> None:
        if not self.enable_oracle_replacement:
            raise ContractPanic("replacement disabled")
        a = self._admin_index()
        if proposition is None:
            self.replacement_propositions[a] = None
            return
        old_index, new_address = proposition
        if not (0 <= old_index < len(self.oracle_addresses)):
            raise ContractPanic("wrong old oracle index")
        if new_address in self.oracle_addresses:
            raise ContractPanic("the oracle is already in the team")
        for i in range(len(self.admins)):                       # reset the votes for this proposition
            self.vote_matrix[i][a] = False
        self.vote_matrix[a][a] = True                           # vote for itself
        self.replacement_propositions[a] = (old_index, new_address)

    def vote_for_a_proposition(self, which_admin: int, support_his_proposition: bool) -> None:
        """contract.cairo:718-738.  The vote is written first and the majority check may then panic
        (`unwrap` of an empty proposition): the transaction reverts, so the written vote is rolled back.
        `which_admin` beyond the admin list is not an error in the contract — the vote matrix is a storage map,
        any key can be written and unknown keys read as false — so such votes are kept in a side table."""
        if not self.enable_oracle_replacement:
            raise ContractPanic("replacement disabled")
        voter = self._admin_index()
        which_admin = int(which_admin)
        if which_admin < 0:
            raise ContractPanic("usize underflow")
        in_range = which_admin < len(self.admins)
        previous = self.vote_matrix[voter][which_admin] if in_range else self._stray_votes.get((voter, which_admin), False)
        self._write_vote(voter, which_admin, bool(support_his_proposition))
        try:
            self._check_for_replacement(which_admin)
        except ContractPanic:
            self._write_vote(voter, which_admin, previous)      # reverted transaction: nothing is stored
            raise

    def _write_vote(self, emitter: int, receiver: int, value: bool) -> None:
        if receiver < len(self.admins):
            self.vote_matrix[emitter][receiver] = value
        else:
            self._stray_votes[(emitter, receiver)] = value

    def _check_for_replacement(self, which: int) -> None:
        a = len(self.admins)
        if which < a:
            votes = sum(1 for i in range(a) if self.vote_matrix[i][which])
        else:
            votes = sum(1 for i in range(a) if self._stray_votes.get((i, which), False))
        if self.required_majority > votes:
            return
        prop = self.replacement_propositions[which] if which < a else None   # storage default: Option::None
        if prop is None:
            raise ContractPanic("Option::unwrap failed")        # `.unwrap()` on an empty proposition
        self.oracle_addresses[prop[0]] = prop[1]
        self._reset_votes()

    def get_replacement_propositions(self):
        if not self.enable_oracle_replacement:
            raise ContractPanic("replacement disabled")
        return list(self.replacement_propositions)

    def get_a_specific_proposition(self, which_admin: int):
        if not self.enable_oracle_replacement:
            raise ContractPanic("replacement disabled")
        return self.replacement_propositions[which_admin]
