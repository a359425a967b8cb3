"""Logger / metric stores / web client / node monitor / CLI (untested in the reference)."""

import datetime
import time

from typer.testing import CliRunner

from p2pfl_b200.cli import app, available_examples
from p2pfl_b200.management.logger import Logger
from p2pfl_b200.management.metric_storage import GlobalMetricStorage, LocalMetricStorage
from p2pfl_b200.management.node_monitor import NodeMonitor
from p2pfl_b200.management.web_services import P2pflWebServices, P2pflWebServicesError
from p2pfl_b200.node_state import NodeState
from p2pfl_b200.settings import Settings


def test_metric_storages():
    loc = LocalMetricStorage()
    loc.add_log("e", 0, "loss", "n1", 1.0, 0)
    loc.add_log("e", 0, "loss", "n1", 0.5, 50)
    loc.add_log("e", 1, "loss", "n2", 0.2, 0)
    assert loc.get_experiment_round_node_logs("e", 0, "n1") == {"loss": [(0, 1.0), (50, 0.5)]}
    assert set(loc.get_experiment_logs("e")) == {0, 1}
    glo = GlobalMetricStorage()
    glo.add_log("e", 0, "acc", "n1", 0.1)
    glo.add_log("e", 0, "acc", "n1", 0.9)  # same round: ignored
    glo.add_log("e", 1, "acc", "n1", 0.5)
    assert glo.get_experiment_node_logs("e", "n1") == {"acc": [(0, 0.1), (1, 0.5)]}


def test_logger_metrics_for_local_and_remote_nodes():
    log = Logger()
    st = NodeState("local-node")
    st.set_experiment("exp", 3)
    log.register_node("local-node", st, False)
    log.log_metric("local-node", "test_loss", 1.5)  # global: uses the node's current round
    log.log_metric("local-node", "train_loss", 0.7, step=10)
    # Q8: a node that lives in another process must not raise
    log.log_metric("10.0.0.9:1234", "test_loss", 2.5, round=0)
    g = log.get_global_logs()["exp"]
    assert g["local-node"]["test_loss"] == [(0, 1.5)] and g["10.0.0.9:1234"]["test_loss"] == [(0, 2.5)]
    assert log.get_local_logs()["exp"][0]["local-node"]["train_loss"] == [(10, 0.7)]
    log.set_level("ERROR")
    assert log.get_level_name(log.get_level()) == "ERROR"
    log.unregister_node("local-node")
    log.cleanup()


class _FakeResp:
    def __init__(self, code=200, body=None):
        self.status_code, self._body, self.text = code, body or {}, "err"

    def json(self):
        return self._body


class _FakeSession:
    def __init__(self):
        self.calls = []

    def post(self, url, json=None, headers=None, timeout=None):
        self.calls.append((url, json, headers))
        if url.endswith("/node"):
            return _FakeResp(200, {"node_id": 7})
        if url.endswith("/fail"):
            return _FakeResp(500)
        return _FakeResp(200)


def test_web_services_client_and_logger_bridge():
    sess = _FakeSession()
    web = P2pflWebServices("http://dash/", "KEY", session=sess)
    web.register_node("n1", True)
    web.send_local_metric("e", 0, "loss", "n1", 0.3, 5)
    web.send_global_metric("e", 0, "acc", "n1", 0.9)
    web.send_system_metric("n1", "cpu", 12.0, datetime.datetime.now())
    web.send_log(datetime.datetime.now(), "n1", 20, "hello")
    paths = [c[0].replace("http://dash", "") for c in sess.calls]
    assert paths == ["/node", "/node-metric/local", "/node-metric/global", "/node-metric/system", "/node-log"]
    assert all(c[2]["x-api-key"] == "KEY" for c in sess.calls)
    assert sess.calls[1][1]["node_id"] == 7
    try:
        web._post("/fail", {})
        raise AssertionError("expected an error")
    except P2pflWebServicesError as e:
        assert e.code == 500
    # the logger forwards metrics and starts a monitor once a dashboard is attached
    Settings.RESOURCE_MONITOR_PERIOD = 0.05
    log = Logger()
    sess2 = _FakeSession()
    log.connect_web("http://dash", "K", session=sess2)
    st = NodeState("n2")
    st.set_experiment("exp", 1)
    log.register_node("n2", st, True)
    log.log_metric("n2", "acc", 0.5)
    time.sleep(0.3)
    log.unregister_node("n2")
    log.cleanup()
    urls = [c[0] for c in sess2.calls]
    assert any(u.endswith("/node-metric/global") for u in urls)
    assert any(u.endswith("/node-metric/system") for u in urls)  # NodeMonitor reported
    Settings.RESOURCE_MONITOR_PERIOD = 1.0


def test_node_monitor_samples():
    seen = []
    mon = NodeMonitor("n", lambda n, k, v: seen.append((n, k, v)))
    sample = mon.sample_host()
    assert "cpu" in sample and "ram" in sample
    assert mon.sample_gpu() == {}  # no GPU index given


def test_cli_lists_and_runs_examples():
    assert {"mnist", "node1", "node2"} <= set(available_examples())
    runner = CliRunner()
    res = runner.invoke(app, ["experiment", "list"])
    assert res.exit_code == 0 and "mnist" in res.stdout
    res = runner.invoke(app, ["login"])
    assert res.exit_code == 0 and "not implemented" in res.stdout
    res = runner.invoke(app, ["experiment", "run", "does-not-exist"])
    assert res.exit_code == 1
    res = runner.invoke(
        app,
        ["experiment", "run", "mnist", "--nodes", "2", "--rounds", "1", "--epochs", "0", "--use_local_protocol",
         "--samples", "640", "--measure_time"],
    )  # fmt: skip
    assert res.exit_code == 0, res.stdout
