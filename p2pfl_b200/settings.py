"""Global run-time configuration.

Same knob names and defaults as the reference's static ``Settings`` class
(reference: p2pfl/settings.py:26-115) so scripts that assign
``Settings.TRAIN_SET_SIZE = 8`` keep working.  Values are read at use time.

Knobs that do not exist in the reference are grouped at the bottom; each one
either switches a documented reference quirk back on or
configures the B200 data plane.
"""

from __future__ import annotations

import os


class Settings:
    """Mutable class-attribute configuration (no instances)."""

    # ---- transport -----------------------------------------------------
    GRPC_TIMEOUT: float = 10.0
    """Per-RPC timeout in seconds."""

    # ---- logging -------------------------------------------------------
    LOG_LEVEL: str = "INFO"
    """DEBUG also makes the learning thread re-raise instead of stopping the node."""
    LOG_DIR: str = "logs"
    """Directory of the optional log file."""
    LOG_TO_FILE: bool = False
    """The reference always writes logs/p2pfl.log; here it is opt-in."""

    # ---- heartbeats ----------------------------------------------------
    HEARTBEAT_PERIOD: float = 2.0
    """Seconds between two beats to the direct neighbours (also the eviction sweep period)."""
    HEARTBEAT_TIMEOUT: float = 5.0
    """A neighbour whose last beat is older than this is dropped."""

    # ---- control-message gossip ---------------------------------------
    GOSSIP_PERIOD: float = 0.1
    """Pacing of the relay thread: at most GOSSIP_MESSAGES_PER_PERIOD sends per period (0 = unpaced)."""
    TTL: int = 10
    """Hops a flooded control message may travel.  It also bounds what a long line of nodes can
    learn about itself (SURVEY Q10): a 64-node line needs ``TTL >= 63``."""
    GOSSIP_MESSAGES_PER_PERIOD: int = 100
    """Send budget of the relay thread per GOSSIP_PERIOD."""
    AMOUNT_LAST_MESSAGES_SAVED: int = 100
    """Size of the ring of message ids used to drop duplicates."""

    # ---- model gossip --------------------------------------------------
    GOSSIP_MODELS_PERIOD: float = 1.0
    """Upper bound on the pause between two model-gossip ticks.  The loop is
    event driven: it wakes as soon as node state changes (SURVEY Appendix C.6)."""
    GOSSIP_MODELS_PER_ROUND: int = 2
    """Peers that get a model per gossip tick."""
    GOSSIP_EXIT_ON_X_EQUAL_ROUNDS: int = 10
    """Model gossip gives up after this many ticks without any change in what the peers hold."""

    # ---- learning ------------------------------------------------------
    TRAIN_SET_SIZE: int = 4
    """Nodes elected to train each round."""
    VOTE_TIMEOUT: float = 60.0
    """Longest wait for the ballots of a round; the tally then uses what arrived."""
    AGGREGATION_TIMEOUT: float = 300.0
    """Longest wait for the models of a round; the aggregate then uses what arrived."""
    WAIT_HEARTBEATS_CONVERGENCE: float = 0.2 * HEARTBEAT_TIMEOUT
    """Pause after start_learning so that beats reach every node before the vote."""

    # ---- monitoring ----------------------------------------------------
    RESOURCE_MONITOR_PERIOD: float = 1.0
    """Sampling period of the node monitor (CPU / RAM / network, GPU through NVML)."""

    # ===== additions (not in the reference) =============================
    VOTE_EVERY_ROUND: bool = True
    """True: re-vote the train set each round (BASELINE config 4).  False
    reproduces the reference's vote-once behaviour (SURVEY Q1)."""
    RESET_OPTIMIZER_EACH_ROUND: bool = True
    """The reference builds a new Trainer (=> new Adam state) per round (Q16)."""
    INIT_MODEL_TIMEOUT: float = 120.0
    """Bounded wait for the initial weights (reference waits forever, Q23)."""
    GRPC_MAX_MESSAGE_BYTES: int = 1 << 30
    """Raised from gRPC's 4 MiB default so >4 MiB models fit (Q9)."""
    GRPC_SERVER_WORKERS: int = 4
    """Threads of a node's gRPC server pool (the reference uses 2)."""
    WIRE_DTYPE: str = "bf16"
    """Wire dtype of the device data plane: 'bf16' (inbox mode B) or 'fp32'."""
    DEVICE_TILE_ELEMS: int = 16384
    """Elements per data-plane tile (one arrival flag per tile)."""
    GIL_SWITCH_INTERVAL: float = float(os.environ.get("P2PFL_GIL_SWITCH_INTERVAL", "0.0005"))
    """``sys.setswitchinterval`` applied when a node starts.  The learning thread shares the
    interpreter with the RPC workers, the gossiper and the heartbeater; with CPython's default
    5 ms a thread returning from a CUDA wait can sit that long before it runs again."""
    OWN_MODEL_PUSH: str = "auto"
    """Exchange policy among trainers when the fused kernel is not used.  "never": the
    reference's policy - send each peer the aggregate of everything it still lacks (partial
    aggregates; fewer messages on sparse overlays, but overlapping partials get rejected and
    re-sent).  "always": every trainer sends only its own model to every train-set peer, which
    can never overlap: exactly |T|(|T|-1) transfers (SURVEY Appendix C, rule 1).  "auto": own
    model on device data planes (any-to-any over NVLink is free), partial aggregates otherwise."""
    PREVOTE: bool = True
    """Cast the ballot for round r+1 while round r trains (ballots are random and
    ``vote_train_set`` already accepts round+1), so the vote is no barrier between rounds."""
    PREVOTE_AHEAD: int = 2
    """How many rounds ahead ballots are cast (with ``PREVOTE``).  One round is enough when every node
    is a hop away; on a long line a ballot needs dozens of relay hops, and the node that finishes a
    round last also votes last, so with 1 every round still ended in a wait for its ballot.  With k,
    a node can be at most k rounds ahead of the slowest one, so ballots arrive for up to round+2k:
    the handler accepts that window."""
    ASYNC_BOOKKEEPING: bool = True
    """``metrics`` / ``models_aggregated`` / ``models_ready`` broadcasts leave through the
    gossiper thread instead of blocking the learning thread on N-1 round trips."""
    BUNDLE_BOOKKEEPING: bool = True
    """Bookkeeping broadcasts that are queued together leave as ONE ``bundle`` message per
    extended neighbour (a reference node gets them one by one): a round's ``models_ready``,
    next round's ``metrics`` and the pre-vote cost one round trip per peer instead of three.
    Every RPC costs both ends ~0.1 ms of interpreter time, which at 8 nodes was the main
    part of the per-round overhead."""
    DIRECT_TRAINSET_STATUS: bool = False
    """``models_aggregated`` (who holds whose model) only matters to the round's trainers.  True: when
    some of them are not direct neighbours it goes to them point to point instead of being flooded
    through every node.  On a 64-node line with 16 trainers that removes the largest share of the
    control messages (35 k -> 6 k per two rounds) and helps the in-memory transport (1.9 -> 1.65 s per
    round), but over gRPC the same run got slower (2.1 -> 4.6 s per round, 8 processes on a CPU-only
    box): the flooded copies ride in bundles between neighbours, the direct ones are 240 extra
    channels and one RPC each.  Off by default."""
    SOCKET_LOCAL_SHORTCUT: bool = True
    """Socket transport: a neighbour that is a node of the SAME process is reached by a direct call into
    its engine (as the in-memory transport does) instead of a frame over the loopback socket.  The
    tests turn it off so that every message of theirs crosses the wire format."""
    DIRECT_ROUND_STATUS: bool = True
    """``models_ready`` (this node holds round r's aggregate) is only ever read for DIRECT neighbours
    (who still needs the model diffused to it: GossipModelStage, reference gossip_model_stage.py:85-106),
    yet the reference floods it through the whole network.  True: it is sent with ttl 1.  On a line of
    64 nodes that is 126 deliveries per round instead of 4032 - the same order as the ballots, which do
    have to reach everybody."""
    BUNDLE_LINGER: float = 0.02
    """Longest a lazily queued broadcast waits for company before it is sent on its own."""
    NVTX_RANGES: bool = os.environ.get("P2PFL_NVTX", "0") not in ("", "0")
    """Wrap every stage of the workflow in an NVTX range (``P2PFL_NVTX=1``), for timeline tools."""
    MLP_EPOCH_KERNEL: bool = True
    """Train the 784-256-128-10 MLP (batch 32) with ``p2pfl_mlp_epoch``: a whole epoch as ONE launch of
    a 16-CTA cluster that keeps the bf16 weights in shared memory, hands activations from layer to
    layer on chip and applies Adam straight out of TMEM.  Other widths / batch sizes, or a device
    that cannot schedule the cluster, use the per-step paths below."""
    MLP_CHAIN_CLUSTER: int = 16
    """Run the fused MLP step as TWO launches: one thread-block cluster of this many CTAs walks
    every GEMM link (forward chain, loss, dgrad chain, then all weight gradients side by side),
    the links separated by the hardware cluster barrier instead of kernel boundaries; then Adam.
    0 = one launch per link (8 per step)."""
    MLP_CHAIN: bool = False
    """Run the dependent links of the fused MLP step (forward chain, loss, dgrad chain) as
    one single-CTA kernel (``p2pfl_gemm_chain``) instead of one launch per link.  Measured
    on B200 (profiles/README.md): 3 launches per step instead of 8, but 13.4 vs 10.6 ms per
    round - one SM streams all the weights (about 0.5 MB per step) and its TMA bandwidth, not
    the launch count, becomes the bound.  Off by default; kept for narrower models."""
    AMP_BF16: bool = False
    """Eager (autograd) learners run forward/backward under bf16 autocast on CUDA."""
    RELAY_PRUNING: bool = True
    """A relay skips neighbours the originator reaches directly (it says so in its
    heartbeats): a full mesh then costs N-1 messages per broadcast, not (N-1)^2."""
    RELAY_PRUNING_MAX_LIST: int = 64
    """A node with more direct neighbours than this does not advertise them (hint size)."""
    FUSED_EXCHANGE: bool = True
    """Use the single-kernel push+FedAvg exchange whenever the whole train set has
    device arenas on distinct GPUs; otherwise models are pulled one by one."""
    FUSED_SNAPSHOT: bool = True
    """The fused exchange keeps a copy of the pre-exchange model (written by the kernel while it
    pushes; one extra fp32 store per element).  An exchange that ends with contributors missing is
    then undone and repeated through add_model messages instead of leaving a partly averaged model."""
    FUSED_INFO_WAIT: float = 30.0
    """Longest a trainer waits for a train-set member's ``arena_info`` before deciding how to
    exchange.  The decision must come out the same on every trainer, so a missing announcement is
    waited for (every node of this framework announces, arena or not), not guessed around."""
    WATCH_SLICE: float = 0.002
    """The native exchange watchdog (``p2pfl_rt_watch``) hands back to Python this often, for the
    liveness checks that need the interpreter (a contributor that left, an experiment that stopped);
    cancellations and the aggregation timeout are handled inside the native wait."""
    ARENA_INFO_WAIT: float = 2.0
    """Longest a sender waits for the announcement of an extended peer that is still building its
    learner before it falls back to a host payload for that peer."""
    DEVICE_INBOX_SLOTS: int = 8
    """Inbox slots per arena = the largest train set the data plane can serve."""
    TRAINER_DIFFUSION_GRACE: float = 0.5
    """Seconds a finished node waits before spending a full-model transfer on a
    train-set peer that has not yet announced ``models_ready``."""
    DEVICE: str = "auto"
    """Default learner device: "auto" (current CUDA device if any), "cpu", "cuda:N"."""
    SAFE_UNPICKLE: bool = True
    """Decode host-path payloads with a numpy-only unpickler (Q21)."""


def set_test_settings() -> None:
    """Shrink every timeout so the suites run in seconds
    (reference: p2pfl/utils.py:37-53)."""
    Settings.GRPC_TIMEOUT = 0.5
    Settings.HEARTBEAT_PERIOD = 0.5
    Settings.HEARTBEAT_TIMEOUT = 2
    Settings.GOSSIP_PERIOD = 0
    Settings.TTL = 10
    Settings.GOSSIP_MESSAGES_PER_PERIOD = 100
    Settings.AMOUNT_LAST_MESSAGES_SAVED = 100
    Settings.GOSSIP_MODELS_PERIOD = 1
    Settings.GOSSIP_MODELS_PER_ROUND = 4
    Settings.GOSSIP_EXIT_ON_X_EQUAL_ROUNDS = 4
    Settings.TRAIN_SET_SIZE = 4
    Settings.VOTE_TIMEOUT = 60
    Settings.AGGREGATION_TIMEOUT = 60
    Settings.WAIT_HEARTBEATS_CONVERGENCE = 0.2 * Settings.HEARTBEAT_TIMEOUT
    Settings.INIT_MODEL_TIMEOUT = 30
    Settings.SOCKET_LOCAL_SHORTCUT = False  # the suites run their nodes in one process: keep them on the wire
    Settings.LOG_LEVEL = "DEBUG"
