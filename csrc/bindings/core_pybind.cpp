// Python bindings of the host runtime (module akka_allreduce_b200._core).
#include <pybind11/functional.h>
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include "../core/checkpoint.h"
#include "../core/config.h"
#include "../core/data_buffer.h"
#include "../core/device_loop.h"
#include "../core/geometry.h"
#include "../core/local_cluster.h"
#include "../core/logging.h"
#include "../core/master.h"
#include "../core/tcp_transport.h"
#include "../core/worker.h"

namespace py = pybind11;
using namespace aar;

namespace {

using FloatArray = py::array_t<float, py::array::c_style | py::array::forcecast>;

// source(iteration) -> 1-D float array ; sink(data, count, iteration)
DataSource wrapSource(py::object fn) {
  return [fn](const AllReduceInputRequest& req) {
    py::gil_scoped_acquire gil;
    FloatArray a = fn(req.iteration).cast<FloatArray>();
    AllReduceInput in;
    in.data.assign(a.data(), a.data() + a.size());
    return in;
  };
}
DataSink wrapSink(py::object fn) {
  return [fn](const AllReduceOutput& out) {
    py::gil_scoped_acquire gil;
    py::array_t<float> d(out.data.size(), out.data.data());
    py::array_t<int> c(out.count.size(), out.count.data());
    fn(d, c, out.iteration);
  };
}

class CallbackPort : public WorkerPort {
 public:
  explicit CallbackPort(py::object fn) : fn_(std::move(fn)) {}
  void send(Address to, Message msg) override {
    py::gil_scoped_acquire gil;
    fn_(to, std::move(msg));
  }

 private:
  py::object fn_;
};

}  // namespace

PYBIND11_MODULE(_core, m) {
  m.doc() = "Host runtime of the B200-native threshold allreduce: protocol engine, coordinator, config, transports";

  py::class_<Geometry>(m, "Geometry")
      .def(py::init<int, int, int>(), py::arg("peers"), py::arg("data_size"), py::arg("max_chunk_size"))
      .def_readonly("P", &Geometry::P)
      .def_readonly("data_size", &Geometry::dataSize)
      .def_readonly("max_chunk_size", &Geometry::maxChunkSize)
      .def_readonly("step_size", &Geometry::stepSize)
      .def("block_start", &Geometry::blockStart)
      .def("block_end", &Geometry::blockEnd)
      .def("block_len", &Geometry::blockLen)
      .def("max_block_len", &Geometry::maxBlockLen)
      .def("num_chunks", &Geometry::numChunks)
      .def("max_num_chunks", &Geometry::maxNumChunks)
      .def("chunk_len", &Geometry::chunkLen)
      .def("total_chunks", &Geometry::totalChunks)
      .def("chunk_index", &Geometry::chunkIndex)
      .def_static("min_required", &Geometry::minRequired)
      .def("min_chunks_required", &Geometry::minChunksRequired);

  py::class_<InitWorkers>(m, "InitWorkers")
      .def(py::init([](std::map<int, Address> workers, Address master, int destId, float thReduce, float thComplete, int maxLag,
                       int dataSize, int maxChunkSize, int generation, int startRound) {
             InitWorkers i;
             i.startRound = startRound;
             i.workers = std::move(workers);
             i.master = master;
             i.destId = destId;
             i.thReduce = thReduce;
             i.thComplete = thComplete;
             i.maxLag = maxLag;
             i.dataSize = dataSize;
             i.maxChunkSize = maxChunkSize;
             i.generation = generation;
             return i;
           }),
           py::arg("workers"), py::arg("master"), py::arg("dest_id"), py::arg("th_reduce"), py::arg("th_complete"), py::arg("max_lag"),
           py::arg("data_size"), py::arg("max_chunk_size"), py::arg("generation") = 0, py::arg("start_round") = 0)
      .def_readwrite("workers", &InitWorkers::workers)
      .def_readwrite("master", &InitWorkers::master)
      .def_readwrite("dest_id", &InitWorkers::destId)
      .def_readwrite("th_reduce", &InitWorkers::thReduce)
      .def_readwrite("th_complete", &InitWorkers::thComplete)
      .def_readwrite("max_lag", &InitWorkers::maxLag)
      .def_readwrite("data_size", &InitWorkers::dataSize)
      .def_readwrite("max_chunk_size", &InitWorkers::maxChunkSize)
      .def_readwrite("generation", &InitWorkers::generation)
      .def_readwrite("start_round", &InitWorkers::startRound);

  py::class_<StartAllreduce>(m, "StartAllreduce")
      .def(py::init([](int round, int generation) { return StartAllreduce{round, generation}; }), py::arg("round"), py::arg("generation") = 0)
      .def_readwrite("round", &StartAllreduce::round)
      .def_readwrite("generation", &StartAllreduce::generation)
      .def("__repr__", [](const StartAllreduce& s) { return "StartAllreduce(" + std::to_string(s.round) + ")"; });

  py::class_<ScatterBlock>(m, "ScatterBlock")
      .def(py::init([](std::vector<float> value, int srcId, int destId, int chunkId, int round) {
             ScatterBlock s;
             s.value = std::move(value);
             s.srcId = srcId;
             s.destId = destId;
             s.chunkId = chunkId;
             s.round = round;
             return s;
           }),
           py::arg("value"), py::arg("src_id"), py::arg("dest_id"), py::arg("chunk_id"), py::arg("round"))
      .def_readwrite("value", &ScatterBlock::value)
      .def_readwrite("src_id", &ScatterBlock::srcId)
      .def_readwrite("dest_id", &ScatterBlock::destId)
      .def_readwrite("chunk_id", &ScatterBlock::chunkId)
      .def_readwrite("round", &ScatterBlock::round)
      .def_readwrite("generation", &ScatterBlock::generation);

  py::class_<ReduceBlock>(m, "ReduceBlock")
      .def(py::init([](std::vector<float> value, int srcId, int destId, int chunkId, int round, int count) {
             ReduceBlock r;
             r.value = std::move(value);
             r.srcId = srcId;
             r.destId = destId;
             r.chunkId = chunkId;
             r.round = round;
             r.count = count;
             return r;
           }),
           py::arg("value"), py::arg("src_id"), py::arg("dest_id"), py::arg("chunk_id"), py::arg("round"), py::arg("count"))
      .def_readwrite("value", &ReduceBlock::value)
      .def_readwrite("src_id", &ReduceBlock::srcId)
      .def_readwrite("dest_id", &ReduceBlock::destId)
      .def_readwrite("chunk_id", &ReduceBlock::chunkId)
      .def_readwrite("round", &ReduceBlock::round)
      .def_readwrite("count", &ReduceBlock::count)
      .def_readwrite("generation", &ReduceBlock::generation);

  py::class_<CompleteAllreduce>(m, "CompleteAllreduce")
      .def(py::init([](int srcId, int round, int generation) { return CompleteAllreduce{srcId, round, generation}; }), py::arg("src_id"),
           py::arg("round"), py::arg("generation") = 0)
      .def_readwrite("src_id", &CompleteAllreduce::srcId)
      .def_readwrite("round", &CompleteAllreduce::round)
      .def_readwrite("generation", &CompleteAllreduce::generation)
      .def("__repr__", [](const CompleteAllreduce& c) {
        return "CompleteAllreduce(" + std::to_string(c.srcId) + ", " + std::to_string(c.round) + ")";
      });

  py::class_<WorkerPort>(m, "WorkerPort");
  py::class_<CallbackPort, WorkerPort>(m, "CallbackPort").def(py::init<py::object>(), py::arg("fn"));

  py::class_<DataBuffer>(m, "DataBuffer")
      .def(py::init<int, int, int, int, std::vector<int>>(), py::arg("row_len"), py::arg("peers"), py::arg("depth"),
           py::arg("max_chunk_size"), py::arg("chunks_per_row"))
      .def("store",
           [](DataBuffer& b, int round, int row, int chunk, std::vector<float> v, int aux) {
             return b.store(round, row, chunk, v.data(), (int)v.size(), aux);
           },
           py::arg("round"), py::arg("row"), py::arg("chunk"), py::arg("value"), py::arg("aux") = 0)
      .def("has", &DataBuffer::has)
      .def("aux", &DataBuffer::aux)
      .def("count_chunk", &DataBuffer::countChunk)
      .def("mask_chunk", &DataBuffer::maskChunk)
      .def("count_round", &DataBuffer::countRound)
      .def("retire", &DataBuffer::retire)
      .def("slot_of", &DataBuffer::slotOf)
      .def("row", [](DataBuffer& b, int round, int p) { return py::array_t<float>(b.rowLen(), b.row(round, p)); })
      .def_property_readonly("depth", &DataBuffer::depth);

  py::class_<WorkerState>(m, "WorkerState")
      .def_readonly("id", &WorkerState::id)
      .def_readonly("th_reduce", &WorkerState::thReduce)
      .def_readonly("th_complete", &WorkerState::thComplete)
      .def_readonly("max_lag", &WorkerState::maxLag)
      .def_readonly("round", &WorkerState::round)
      .def_readonly("max_round", &WorkerState::maxRound)
      .def_readonly("max_scattered", &WorkerState::maxScattered)
      .def_readonly("completed", &WorkerState::completed)
      .def_readonly("data_size", &WorkerState::dataSize)
      .def_readonly("max_chunk_size", &WorkerState::maxChunkSize)
      .def_readonly("num_peers", &WorkerState::numPeers)
      .def_readonly("generation", &WorkerState::generation);

  py::class_<Worker>(m, "Worker")
      .def(py::init([](py::object source, py::object sink, WorkerPort* port) {
             return std::make_unique<Worker>(wrapSource(source), wrapSink(sink), port);
           }),
           py::arg("source"), py::arg("sink"), py::arg("port"), py::keep_alive<1, 4>())
      .def("receive", [](Worker& w, const Message& msg) { w.receive(msg); })
      .def_property_readonly("id", &Worker::id)
      .def_property_readonly("round", &Worker::round)
      .def_property_readonly("max_round", &Worker::maxRound)
      .def_property_readonly("max_scattered", &Worker::maxScattered)
      .def_property_readonly("completed", [](const Worker& w) { return std::vector<int>(w.completed().begin(), w.completed().end()); })
      .def_property_readonly("initialized", &Worker::initialized)
      .def_property_readonly("geometry", &Worker::geometry)
      .def("state", &Worker::state)
      .def("metrics", [](const Worker& w) { return w.metrics().asMap(); })
      .def("contributor_histogram", [](const Worker& w) {
        return std::vector<int64_t>(w.metrics().contributorHist.begin(), w.metrics().contributorHist.end());
      });

  // ---- device-plane round loop (csrc/core/device_loop.h): Python is the caller's source and sink, nothing else per round
  py::class_<DevicePlane>(m, "DevicePlane")
      .def(py::init([](uintptr_t engine, uintptr_t enqueue, uintptr_t doorbell, uintptr_t flushed, uintptr_t synchronize, int slots, int maxLag,
                       long long dataSize, unsigned extraFlags) {
             DevicePlane p;
             p.engine = reinterpret_cast<void*>(engine);
             p.enqueue = reinterpret_cast<DevicePlane::EnqueueFn>(enqueue);
             p.doorbell = reinterpret_cast<DevicePlane::DoorbellFn>(doorbell);
             p.flushed = reinterpret_cast<DevicePlane::FlushedFn>(flushed);
             p.synchronize = reinterpret_cast<DevicePlane::SyncFn>(synchronize);
             p.slots = slots;
             p.maxLag = maxLag;
             p.dataSize = dataSize;
             p.extraFlags = extraFlags;
             return p;
           }),
           py::arg("engine"), py::arg("enqueue"), py::arg("doorbell"), py::arg("flushed"), py::arg("synchronize") = 0, py::arg("slots") = 2,
           py::arg("max_lag") = 1, py::arg("data_size") = 0, py::arg("extra_flags") = 0u)
      .def_readonly("slots", &DevicePlane::slots)
      .def_readonly("max_lag", &DevicePlane::maxLag)
      .def_readonly("data_size", &DevicePlane::dataSize);

  py::class_<DeviceLoop>(m, "DeviceLoop")
      .def(py::init([](py::object source, py::object sink, WorkerPort* port) {
             // source(round) -> (input ptr, output ptr, counts ptr, stream, numel) ; sink(round, slot)
             return std::make_unique<DeviceLoop>(
                 [source](int round) {
                   py::gil_scoped_acquire gil;
                   py::tuple t = source(round).cast<py::tuple>();
                   RoundBuffers b;
                   b.input = t[0].cast<uintptr_t>();
                   b.output = t[1].cast<uintptr_t>();
                   b.counts = t[2].cast<uintptr_t>();
                   b.stream = t[3].cast<uintptr_t>();
                   b.numel = t[4].cast<long long>();
                   return b;
                 },
                 [sink](int round, int slot) {
                   py::gil_scoped_acquire gil;
                   sink(round, slot);
                 },
                 port);
           }),
           py::arg("source"), py::arg("sink"), py::arg("port"), py::keep_alive<1, 4>())
      .def("set_init",
           [](DeviceLoop& l, py::object fn) {
             l.setInit([fn](const InitWorkers& msg) {
               py::gil_scoped_acquire gil;
               fn(msg);
             });
           })
      .def("set_board",
           [](DeviceLoop& l, py::object post, py::object newest) {
             l.setBoard(
                 [post](int r) {
                   py::gil_scoped_acquire gil;
                   post(r);
                 },
                 [newest]() {
                   py::gil_scoped_acquire gil;
                   return newest().cast<int>();
                 });
           })
      .def("clear_board", [](DeviceLoop& l) { l.setBoard(nullptr, nullptr); })
      .def("set_plane", &DeviceLoop::setPlane, py::arg("plane"), py::arg("id"), py::arg("generation") = 0, py::arg("master") = kNoAddress,
           py::arg("start_round") = 0)
      .def("clear_plane", &DeviceLoop::clearPlane)
      .def("set_drop_start", &DeviceLoop::setDropStart, py::arg("drop"))
      .def("receive", [](DeviceLoop& l, const Message& msg) { l.receive(msg); }, py::call_guard<py::gil_scoped_release>())
      .def("on_start", &DeviceLoop::onStart, py::arg("round"), py::call_guard<py::gil_scoped_release>())
      .def("peer_ahead", &DeviceLoop::peerAhead, py::call_guard<py::gil_scoped_release>())
      .def("deliver_flushed", &DeviceLoop::deliverFlushed, py::arg("block") = false, py::call_guard<py::gil_scoped_release>())
      .def("drain", &DeviceLoop::drain, py::call_guard<py::gil_scoped_release>())
      .def("run", &DeviceLoop::run, py::arg("node"), py::arg("timeout_sec") = 3600.0, py::call_guard<py::gil_scoped_release>())
      .def_property_readonly("busy", &DeviceLoop::busy)
      .def_property_readonly("id", &DeviceLoop::id)
      .def_property_readonly("generation", &DeviceLoop::generation)
      .def_property_readonly("round", &DeviceLoop::round)
      .def_property_readonly("max_round", &DeviceLoop::maxRound)
      .def_property_readonly("max_scattered", &DeviceLoop::maxScattered)
      .def_property_readonly("start_round", &DeviceLoop::startRound)
      .def_property_readonly("completed",
                             [](const DeviceLoop& l) {
                               const std::set<int> c = l.completed();
                               return std::vector<int>(c.begin(), c.end());
                             })
      .def("stats", [](const DeviceLoop& l) {
        const DeviceLoop::Snapshot snap = l.snapshot();   // one consistent view, safe while the loop runs in another thread
        const DeviceLoopStats& s = snap.stats;
        py::dict d;
        d["round"] = snap.round;
        d["max_round"] = snap.maxRound;
        d["max_scattered"] = snap.maxScattered;
        d["completed"] = std::vector<int>(snap.completed.begin(), snap.completed.end());
        d["rounds_delivered"] = s.delivered;
        d["rounds_out_of_order"] = s.outOfOrder;
        d["self_started_rounds"] = s.selfStarts;
        d["rounds_forced"] = s.forced;
        d["rounds_launched"] = s.launched;
        py::dict h;
        for (auto& kv : s.lagHist) h[py::int_(kv.first)] = kv.second;
        d["lag_hist"] = h;
        return d;
      });

  py::class_<MasterConfig>(m, "MasterConfig")
      .def(py::init([](int totalWorkers, float thAllreduce, float thReduce, float thComplete, int maxLag, int dataSize, int maxRound,
                       int maxChunkSize, int pipelineDepth, int startRound) {
             MasterConfig c;
             c.startRound = startRound;
             c.totalWorkers = totalWorkers;
             c.thAllreduce = thAllreduce;
             c.thReduce = thReduce;
             c.thComplete = thComplete;
             c.maxLag = maxLag;
             c.dataSize = dataSize;
             c.maxRound = maxRound;
             c.maxChunkSize = maxChunkSize;
             c.pipelineDepth = pipelineDepth;
             return c;
           }),
           py::arg("total_workers") = 2, py::arg("th_allreduce") = 1.f, py::arg("th_reduce") = 1.f, py::arg("th_complete") = 1.f,
           py::arg("max_lag") = 1, py::arg("data_size") = 10, py::arg("max_round") = 100, py::arg("max_chunk_size") = 2,
           py::arg("pipeline_depth") = 0, py::arg("start_round") = 0)
      .def_readwrite("total_workers", &MasterConfig::totalWorkers)
      .def_readwrite("th_allreduce", &MasterConfig::thAllreduce)
      .def_readwrite("th_reduce", &MasterConfig::thReduce)
      .def_readwrite("th_complete", &MasterConfig::thComplete)
      .def_readwrite("max_lag", &MasterConfig::maxLag)
      .def_readwrite("data_size", &MasterConfig::dataSize)
      .def_readwrite("max_round", &MasterConfig::maxRound)
      .def_readwrite("max_chunk_size", &MasterConfig::maxChunkSize)
      .def_readwrite("pipeline_depth", &MasterConfig::pipelineDepth)
      .def_readwrite("start_round", &MasterConfig::startRound);

  py::class_<Master>(m, "Master")
      .def(py::init<MasterConfig, WorkerPort*, Address>(), py::arg("config"), py::arg("port"), py::arg("self_address") = kNoAddress,
           py::keep_alive<1, 3>())
      .def("member_up", &Master::memberUp, py::arg("address"), py::arg("role") = "worker")
      .def("terminated", &Master::terminated)
      .def("receive", [](Master& ms, const Message& msg) { ms.receive(msg); })
      .def_property_readonly("round", &Master::round)
      .def_property_readonly("quorum_round", &Master::quorumRound)
      .def_property_readonly("num_complete", &Master::numComplete)
      .def_property_readonly("generation", &Master::generation)
      .def_property_readonly("started", &Master::started)
      .def_property_readonly("finished", &Master::finished)
      .def_property_readonly("workers", &Master::workers)
      .def("metrics", [](const Master& ms) { return ms.metrics().asMap(); });

  py::class_<LocalCluster>(m, "LocalCluster")
      .def(py::init([](MasterConfig cfg, int n, py::object makeSource, py::object makeSink) {
             return std::make_unique<LocalCluster>(
                 cfg, n, [makeSource](int i) { return wrapSource(makeSource(i)); }, [makeSink](int i) { return wrapSink(makeSink(i)); });
           }),
           py::arg("config"), py::arg("num_workers"), py::arg("make_source"), py::arg("make_sink"))
      .def("join_all", &LocalCluster::joinAll)
      .def("join", &LocalCluster::join)
      .def("run", &LocalCluster::run, py::arg("max_messages") = -1)
      .def("hold", &LocalCluster::hold)
      .def("release", &LocalCluster::release)
      .def("terminate", &LocalCluster::terminate)
      .def("set_drop",
           [](LocalCluster& c, py::object fn) {
             if (fn.is_none()) {
               c.setDrop(nullptr);
             } else {
               c.setDrop([fn](Address from, Address to, const Message& msg) { return fn(from, to, msg).cast<bool>(); });
             }
           })
      .def("master", &LocalCluster::master, py::return_value_policy::reference_internal)
      .def("worker", &LocalCluster::worker, py::return_value_policy::reference_internal)
      .def_property_readonly("num_workers", &LocalCluster::numWorkers)
      .def_property_readonly("delivered", &LocalCluster::delivered)
      .def_property_readonly("dropped", &LocalCluster::dropped)
      .def_property_readonly("pending", &LocalCluster::pending)
      .def_static("worker_addr", &LocalCluster::workerAddr)
      .def_property_readonly_static("MASTER_ADDR", [](py::object) { return LocalCluster::kMasterAddr; });

  m.def(
      "cpu_baseline_seconds",
      [](const MasterConfig& cfg, int n) {
        double checksum = 0;
        int64_t msgs = 0;
        double sec;
        {
          py::gil_scoped_release rel;
          sec = cpuBaselineSeconds(cfg, n, &checksum, &msgs);
        }
        return py::make_tuple(sec, checksum, msgs);
      },
      py::arg("config"), py::arg("num_workers"));

  py::class_<Config>(m, "Config")
      .def_static("parse_string", &Config::parseString)
      .def_static("parse_file", &Config::parseFile)
      .def("with_fallback", &Config::withFallback)
      .def("has", &Config::has)
      .def("get_string", py::overload_cast<const std::string&>(&Config::getString, py::const_))
      .def("get_int", py::overload_cast<const std::string&>(&Config::getInt, py::const_))
      .def("get_double", py::overload_cast<const std::string&>(&Config::getDouble, py::const_))
      .def("get_bool", py::overload_cast<const std::string&>(&Config::getBool, py::const_))
      .def("get_list", &Config::getList)
      .def("get_duration_ms", py::overload_cast<const std::string&>(&Config::getDurationMs, py::const_))
      .def("set", &Config::set)
      .def("set_list", &Config::setList)
      .def("keys", &Config::keys)
      .def("render", &Config::render);

  py::class_<AllreduceSettings>(m, "AllreduceSettings")
      .def_static("from_config", &AllreduceSettings::fromConfig)
      .def_readwrite("master", &AllreduceSettings::master)
      .def_readwrite("dtype", &AllreduceSettings::dtype)
      .def_readwrite("algo", &AllreduceSettings::algo)
      .def_readwrite("normalize", &AllreduceSettings::normalize)
      .def_readwrite("quorum_policy", &AllreduceSettings::quorumPolicy)
      .def_readwrite("spin_timeout_us", &AllreduceSettings::spinTimeoutUs)
      .def_readwrite("hostname", &AllreduceSettings::hostname)
      .def_readwrite("port", &AllreduceSettings::port)
      .def_readwrite("heartbeat_ms", &AllreduceSettings::heartbeatMs)
      .def_readwrite("auto_down_after_ms", &AllreduceSettings::autoDownAfterMs)
      .def_readwrite("loglevel", &AllreduceSettings::loglevel);

  m.def("save_worker_checkpoint", &saveWorkerCheckpoint);
  m.def("load_worker_checkpoint", &loadWorkerCheckpoint);
  m.def("save_master_checkpoint", &saveMasterCheckpoint);
  m.def("load_master_checkpoint", &loadMasterCheckpoint);
  m.def("resume_master_from_checkpoint", &resumeMasterFromCheckpoint);

  py::class_<RingSnapshot>(m, "RingSnapshot")
      .def(py::init<>())
      .def_readwrite("rank", &RingSnapshot::rank)
      .def_readwrite("P", &RingSnapshot::P)
      .def_readwrite("L", &RingSnapshot::L)
      .def_readwrite("data_size", &RingSnapshot::dataSize)
      .def_readwrite("max_chunk_size", &RingSnapshot::maxChunkSize)
      .def_readwrite("round", &RingSnapshot::round)
      .def_readwrite("scatter_ring", &RingSnapshot::scatterRing)
      .def_readwrite("reduce_ring", &RingSnapshot::reduceRing)
      .def_readwrite("scatter_flags", &RingSnapshot::scatterFlags)
      .def_readwrite("reduce_flags", &RingSnapshot::reduceFlags);
  m.def("save_ring_snapshot", &saveRingSnapshot);
  m.def("load_ring_snapshot", &loadRingSnapshot);

  py::class_<MasterNode>(m, "MasterNode")
      .def(py::init<MasterConfig, const std::string&, int, double>(), py::arg("config"), py::arg("host") = "127.0.0.1", py::arg("port") = 0,
           py::arg("auto_down_after_ms") = 10000.0)
      .def_property_readonly("port", &MasterNode::port)
      .def("master", &MasterNode::master, py::return_value_policy::reference_internal)
      .def("poll", &MasterNode::poll, py::arg("timeout_ms") = 0, py::call_guard<py::gil_scoped_release>())
      .def("run", &MasterNode::run, py::arg("timeout_sec") = 60.0, py::call_guard<py::gil_scoped_release>())
      .def("broadcast_shutdown", &MasterNode::broadcastShutdown, py::call_guard<py::gil_scoped_release>())
      .def("set_checkpoint", &MasterNode::setCheckpoint, py::arg("path"), py::arg("every_rounds") = 100)
      .def("checkpoint_now", &MasterNode::checkpointNow);

  py::class_<WorkerNode, WorkerPort>(m, "WorkerNode")
      .def(py::init<const std::string&, int, const std::string&, int, double, const std::string&>(), py::arg("master_host"),
           py::arg("master_port"), py::arg("listen_host") = "127.0.0.1", py::arg("listen_port") = 0, py::arg("heartbeat_ms") = 1000.0,
           py::arg("role") = "worker")
      .def("set_handler",
           [](WorkerNode& n, py::object fn) {
             n.setHandler([fn](const Message& msg) {
               py::gil_scoped_acquire gil;
               fn(msg);
             });
           })
      // Route inbound messages straight into a C++ worker (no Python on the data path).
      .def("attach", [](WorkerNode& n, Worker& w) { n.setHandler([&w](const Message& msg) { w.receive(msg); }); }, py::keep_alive<1, 2>())
      .def_property_readonly("listen_port", &WorkerNode::listenPort)
      .def_property_readonly("shutdown_requested", &WorkerNode::shutdownRequested)
      .def_property_readonly("master_alive", &WorkerNode::masterAlive)
      .def_property_readonly("shutdown_by_master", &WorkerNode::shutdownByMaster)
      .def("poll", &WorkerNode::poll, py::arg("timeout_ms") = 0, py::call_guard<py::gil_scoped_release>())
      .def("run", &WorkerNode::run, py::arg("timeout_sec") = 60.0, py::call_guard<py::gil_scoped_release>())
      .def("send_complete",
           [](WorkerNode& n, int srcId, int round, int generation) { n.send(WorkerNode::kMaster, CompleteAllreduce{srcId, round, generation}); })
      .def_property_readonly_static("MASTER", [](py::object) { return WorkerNode::kMaster; });

  m.def("set_log_level", [](const std::string& l) { Logger::instance().setLevel(Logger::parse(l)); });
}
