"""L1: ClusterSpec / Server / devices (Hogwild/Hogwild.py:20-33,
Basics-Tutorial/Servers.ipynb)."""
import socket

import pytest
import torch


def test_cluster_spec_dict_and_indexing(dtg):
    cs = dtg.ClusterSpec({"ps": ["localhost:2222"],
                          "worker": ["localhost:2223", "localhost:2224"]})
    assert cs.jobs == ["ps", "worker"]
    assert cs.num_tasks("worker") == 2 and cs.num_tasks("ps") == 1
    assert cs.task_address("worker", 1) == "localhost:2224"
    assert cs.task_indices("worker") == [0, 1]
    assert cs.as_dict()["worker"] == ["localhost:2223", "localhost:2224"]
    with pytest.raises(ValueError):
        cs.task_address("worker", 2)
    with pytest.raises(ValueError):
        cs.num_tasks("evaluator")
    # ps first in the dense rank -> GPU map
    assert cs.global_rank("ps", 0) == 0 and cs.global_rank("worker", 1) == 2


def test_cluster_spec_short_form_uses_reference_ports(dtg):
    cs = dtg.ClusterSpec("ps=1,worker=2")
    assert cs.as_dict() == {"ps": ["localhost:2222"],
                            "worker": ["localhost:2223", "localhost:2224"]}
    assert 'name: "ps"' in cs.as_cluster_def()


def test_sparse_task_dict(dtg):
    cs = dtg.ClusterSpec({"worker": {1: "h:1", 3: "h:3"}})
    assert cs.task_indices("worker") == [1, 3]
    assert cs.as_dict() == {"worker": {1: "h:1", 3: "h:3"}}


def test_device_spec_merge_and_canonical(dtg):
    from dtg_b200.cluster import DeviceSpec
    outer = DeviceSpec.from_string("/job:worker/task:1")
    inner = DeviceSpec.from_string("/job:ps/task:0")
    assert outer.merge_from(inner).to_string() == "/job:ps/task:0"
    assert DeviceSpec.from_string("/cpu:0").device_type == "CPU"
    assert DeviceSpec.from_string("/gpu:0").device_type == "GPU"
    assert dtg.canonical_device("/job:ps/task:0") == "/job:ps/replica:0/task:0/device:CPU:0"
    with pytest.raises(ValueError):
        DeviceSpec.from_string("job:ps")


def test_server_target_and_port_listening(dtg, ports):
    (p,) = ports(1)
    server = dtg.Server({"worker": ["localhost:%d" % p]}, job_name="worker", task_index=0)
    try:
        # Servers.ipynb:80 -- 'grpc://localhost:2222'
        assert server.target == "grpc://localhost:%d" % p
        # Servers.ipynb:112-120 -- lsof shows the port listening
        with socket.create_connection(("127.0.0.1", p), timeout=2):
            pass
        assert 'protocol: "grpc"' in server.server_def
        assert "task_index: 0" in server.server_def
    finally:
        server.stop()


def test_create_local_server_and_list_devices(dtg):
    server = dtg.Server.create_local_server()
    try:
        assert server.target.startswith("grpc://localhost:")
        assert not server.target.endswith(":0")          # ephemeral port got resolved
        with dtg.Session(server.target) as sess:
            devs = sess.list_devices()
        assert devs[0].name == "/job:local/replica:0/task:0/device:CPU:0"
        assert devs[0].device_type == "CPU"
    finally:
        server.stop()


def test_remote_variable_ops_execute_on_owner(dtg, ports):
    """A 'ps' server and a 'worker' server in one process: the worker's ops on a
    ps-placed variable travel over the wire and mutate the ps host's state."""
    p0, p1 = ports(2)
    spec = {"ps": ["localhost:%d" % p0], "worker": ["localhost:%d" % p1]}
    ps = dtg.Server(spec, job_name="ps", task_index=0)
    wk = dtg.Server(spec, job_name="worker", task_index=0)   # becomes current runtime
    try:
        with dtg.device("/job:ps/task:0"):
            v = dtg.Variable(torch.tensor([1.0, 2.0]), name="w")
        v.initializer.run()
        assert torch.equal(ps.host.vars["w"], torch.tensor([1.0, 2.0]))
        opt = dtg.train.GradientDescentOptimizer(0.5)
        opt.apply_gradients([(torch.tensor([2.0, 2.0]), v)]).run()
        assert torch.equal(ps.host.vars["w"], torch.tensor([0.0, 1.0]))
        assert torch.equal(v.read_value(), torch.tensor([0.0, 1.0]))
        assert "w" not in wk.host.vars
    finally:
        wk.stop()
        ps.stop()


def test_ps_join_returns_when_workers_done(dtg, ports):
    import threading
    (p0, p1) = ports(2)
    spec = {"ps": ["localhost:%d" % p0], "worker": ["localhost:%d" % p1]}
    ps = dtg.Server(spec, job_name="ps", task_index=0)
    t = threading.Thread(target=ps.join, daemon=True)
    t.start()
    from dtg_b200.parallel.rpc import HostClient
    c = HostClient("localhost", p0)
    assert c.ping() == "ps:0"
    c.mark_done("worker:0")
    t.join(timeout=5)
    assert not t.is_alive()
    c.close()


def test_unavailable_error_is_a_recoverable_aborted_error(dtg, ports):
    """README.md:78: MonitoredSession recovers from Aborted / Unavailable; a ps that is not
    listening raises the latter, which every ``except AbortedError`` path also catches."""
    from dtg_b200.parallel.rpc import HostClient
    assert issubclass(dtg.errors.UnavailableError, dtg.errors.AbortedError)
    c = HostClient("localhost", ports(1)[0], connect_timeout=0.2)
    with pytest.raises(dtg.errors.UnavailableError):
        c.num_done()
    with pytest.raises(dtg.errors.AbortedError):
        c.num_done()
