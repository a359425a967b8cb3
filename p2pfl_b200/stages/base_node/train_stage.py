"""Stage 3a (trainers): evaluate, train, contribute, exchange partial aggregates
(reference: p2pfl/stages/base_node/train_stage.py:42-184).

Exchange policy: a trainer keeps pushing, to every train-set peer that still
lacks something it holds, the aggregate of exactly the entries that peer does
not have (``models_aggregated`` announcements tell who has what).  The loop is
woken by those announcements instead of sleeping a period.
"""

from __future__ import annotations

import time
from typing import Any, List, Optional, Type

from p2pfl_b200.management import tracer
from p2pfl_b200.management.logger import logger
from p2pfl_b200.stages.common import announce_aggregated, broadcast_metrics, evaluate_and_share, pack_model
from p2pfl_b200.stages.stage import LearningContext, Stage
from p2pfl_b200.stages.stage_factory import StageFactory


class TrainStage(Stage):
    @staticmethod
    def name() -> str:
        return "TrainStage"

    @staticmethod
    def execute(ctx: LearningContext) -> Optional[Type[Stage]]:
        st, proto, agg = ctx.state, ctx.communication_protocol, ctx.aggregator
        if st.learner is None or st.round is None:
            return None
        t0 = time.time()
        agg.set_nodes_to_aggregate(st.train_set)
        TrainStage.replay_early_models(ctx)
        st.learner.begin_round(st.round)
        ctx.timings["arm"] = ctx.timings.get("arm", 0.0) + time.time() - t0

        if ctx.early_stopping_fn():
            return None
        # How this round's models will be exchanged is decided BEFORE any device work is queued (it
        # depends on the vote and on the arenas the members announced, not on the training).
        fused = None
        plan = getattr(proto, "plan_fused_exchange", None)
        if plan is not None and getattr(st.learner, "arena", None) is not None:
            fused = plan(st, st.learner)
        seq = st.seq_base + st.round + 1
        will_prevote = TrainStage.will_prevote(ctx)
        pipelined = fused is not None and fused.can_prelaunch and hasattr(st.learner, "evaluate_async")
        handle = st.learner.evaluate_async() if pipelined else None
        result = None
        if handle is not None:
            # Device pipeline of the round: evaluation, the epoch(s), validation and the exchange are
            # queued back to back; the host reads their results afterwards.  Between the last training
            # kernel and the exchange kernel there is no host round trip any more.
            t1 = time.time()
            logger.info(st.addr, "training")
            launched = []

            def queued() -> None:
                fused.launch(seq)
                launched.append(1)
                # The pre-vote (and what waits in its bundle) leaves NOW, when the whole round is
                # queued on the device and this thread is about to block on its results: every node
                # broadcasts at this point of its round, so the seven inbound bundles are handled while
                # the hosts wait - not while they issue the launches (handlers and launches share the
                # interpreter: at 8 nodes the launch phase took 0.6 ms longer than alone).
                if will_prevote:
                    TrainStage.prevote(ctx)

            st.learner.fit(after_launch=queued)
            if will_prevote:
                TrainStage.prevote(ctx)  # no-op when cast above (fit may not have got that far)
            t2 = time.time()
            results = st.learner.evaluate_finish(handle)
            broadcast_metrics(proto, st, results, lazy=True)  # nobody waits on metrics
            if launched:
                result = fused.finish()  # always collected: its watchdog is what stops a spinning kernel
            if ctx.early_stopping_fn():
                return None
            if not launched:
                result = fused.run(seq)
        else:
            # the metrics wait (lazily queued) for the pre-vote that follows the evaluation: the two
            # - and last round's models_ready - reach each neighbour as one bundled message
            evaluate_and_share(ctx, lazy=will_prevote)
            if will_prevote:
                TrainStage.prevote(ctx)
            t1 = time.time()

            if ctx.early_stopping_fn():
                return None
            logger.info(st.addr, "training")
            st.learner.fit()
            t2 = time.time()

            if ctx.early_stopping_fn():
                return None
            if fused is not None:
                # device data plane: ONE kernel pushes this model to every trainer and
                # FedAvg-accumulates theirs into the live parameters as the tiles land
                result = fused.run(seq)
        if result is not None:
            contributors, total_w = result
            held = agg.add_model(st.learner.live_params(), contributors, total_w)
            # every trainer that ran the exchange to completion holds the aggregate by
            # construction: GossipModelStage need not wait for their `models_ready`
            st.fused_complete_round = st.round
        else:
            # no fused exchange this round, or it did not complete (the local model is back in
            # place and everybody was told): contribute through add_model messages
            st.fused_complete_round = None
            held = agg.add_model(st.learner.export_model(), [st.addr], st.learner.get_num_samples()[0])
            announce_aggregated(proto, st, held)
            TrainStage._gossip_partial_aggregates(ctx)
        t3 = time.time()
        for k, dt in (("evaluate", t1 - t0), ("fit", t2 - t1), ("exchange", t3 - t2)):
            ctx.timings[k] = ctx.timings.get(k, 0.0) + dt
        if tracer.enabled():
            how = "fused" if result is not None else "messages"
            tracer.complete("evaluate" if handle is None else "arm + evaluate launch", st.addr, t0, t1, "phase", {"round": st.round})
            tracer.complete("fit", st.addr, t1, t2, "phase", {"round": st.round, "pipelined": handle is not None})
            tracer.complete("exchange", st.addr, t2, t3, "phase", {"round": st.round, "how": how})
        if ctx.early_stopping_fn():
            return None
        return StageFactory.get_stage("GossipModelStage")

    @staticmethod
    def will_prevote(ctx: LearningContext) -> bool:
        from p2pfl_b200.settings import Settings

        st = ctx.state
        if not (Settings.PREVOTE and Settings.VOTE_EVERY_ROUND):
            return False
        return not (st.round is None or st.total_rounds is None or st.round + 1 >= st.total_rounds)

    @staticmethod
    def prevote(ctx: LearningContext) -> None:
        """Cast the ballots of the next ``PREVOTE_AHEAD`` rounds now (those not cast yet), so that they
        travel while this round trains."""
        from p2pfl_b200.settings import Settings

        st = ctx.state
        if not TrainStage.will_prevote(ctx):
            return
        from p2pfl_b200.stages.base_node.vote_train_set_stage import VoteTrainSetStage

        # further than one round ahead only among nodes of this framework: a reference node files
        # ballots for its current and next round and drops later ones (vote_train_set_command.py:54-67)
        proto = ctx.communication_protocol
        extended = getattr(proto, "_extended", None)
        known = proto.get_neighbors(only_direct=False)
        ours_only = extended is not None and all(n in extended for n in known)
        for ahead in range(1, (max(1, Settings.PREVOTE_AHEAD) if ours_only else 1) + 1):
            if st.round is None or st.total_rounds is None or st.round + ahead >= st.total_rounds:
                break
            VoteTrainSetStage.cast_ballot(ctx, st.round + ahead, background=True)

    @staticmethod
    def replay_early_models(ctx: LearningContext) -> None:
        """Models that arrived before the vote finished were stashed by
        AddModelCommand; hand them to the aggregator now."""
        st = ctx.state
        with st.stash_lock:
            stash, st.pending_models = st.pending_models, []
        for rnd, thunk in stash:
            if rnd == st.round:
                thunk()

    @staticmethod
    def _gossip_partial_aggregates(ctx: LearningContext) -> None:
        st, proto, agg = ctx.state, ctx.communication_protocol, ctx.aggregator

        from p2pfl_b200.settings import Settings

        own_only = Settings.OWN_MODEL_PUSH == "always" or (
            Settings.OWN_MODEL_PUSH == "auto" and getattr(proto, "device_payloads", False)
        )

        def has(n: str) -> List[str]:
            return st.models_aggregated.get(n, [])

        def candidates() -> List[str]:
            mine = {st.addr} if own_only else set(agg.get_aggregated_models())
            known = proto.get_neighbors(only_direct=False)
            return [n for n in st.train_set if n != st.addr and n in known and (mine - set(has(n)))]

        def status() -> Any:
            return sorted((n, tuple(sorted(has(n)))) for n in st.train_set if n != st.addr)

        cache = {}

        def model_fn(n: str) -> Any:
            if st.learner is None or st.round is None:
                return None
            if own_only:
                # everything but my own entry is "already there" from this sender's point of view
                others = [c for c in agg.get_aggregated_models() if c != st.addr]
                model, contributors, weight = agg.get_partial_aggregation(others + has(n))
            else:
                model, contributors, weight = agg.get_partial_aggregation(has(n))
            if model is None:
                return None
            key = tuple(sorted(contributors))
            if key not in cache:  # same partial for several peers: pack once (Q13)
                cache[key] = pack_model(proto, st.learner, model)
            return proto.build_weights("add_model", st.round, cache[key], contributors, weight)

        def changed(snapshot: Any) -> bool:
            return ctx.early_stopping_fn() or status() != snapshot or not candidates()

        def wait_fn(t: float) -> None:
            snap = status()
            st.wait_for(lambda: changed(snap), t)

        proto.gossip_weights(ctx.early_stopping_fn, candidates, status, model_fn, create_connection=True, wait_fn=wait_fn)
