// NUMA / CPU-affinity binding for spawned ranks.
//
// hfai.multiprocessing.spawn(..., bind_numa=True) (ref:restnet_ddp.py:155) pins each child to
// the NUMA node of its GPU; hfai is closed, so the behaviour is re-derived: PCI bus id -> sysfs
// numa_node -> node cpulist -> sched_setaffinity.  Keeping the data-loader threads and the launch
// thread on the GPU's socket avoids cross-socket pinned-memory traffic on the H2D path.
#ifndef _GNU_SOURCE
#define _GNU_SOURCE
#endif
#include "pdb200_runtime.h"

#include <sched.h>

#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

static bool read_file(const std::string& path, std::string* out) {
  FILE* f = std::fopen(path.c_str(), "r");
  if (!f) return false;
  char buf[4096];
  size_t n = std::fread(buf, 1, sizeof(buf) - 1, f);
  std::fclose(f);
  buf[n] = 0;
  *out = buf;
  return true;
}

// "0-15,32-47" -> cpu_set_t
static int parse_cpulist(const std::string& s, cpu_set_t* set) {
  CPU_ZERO(set);
  int count = 0;
  const char* p = s.c_str();
  while (*p) {
    while (*p && !std::isdigit(static_cast<unsigned char>(*p))) ++p;
    if (!*p) break;
    char* end = nullptr;
    long a = std::strtol(p, &end, 10), b = a;
    p = end;
    if (*p == '-') {
      b = std::strtol(p + 1, &end, 10);
      p = end;
    }
    for (long c = a; c <= b && c < CPU_SETSIZE; ++c) {
      CPU_SET(static_cast<int>(c), set);
      ++count;
    }
  }
  return count;
}

extern "C" int pdb_bind_numa_node(int node) {
  if (node < 0) return -1;
  std::string list;
  if (!read_file("/sys/devices/system/node/node" + std::to_string(node) + "/cpulist", &list)) return -1;
  cpu_set_t want, have, both;
  if (parse_cpulist(list, &want) == 0) return -1;
  // Intersect with the current mask: a container cgroup may expose only a subset.
  if (sched_getaffinity(0, sizeof(have), &have) != 0) return -2;
  CPU_AND(&both, &want, &have);
  if (CPU_COUNT(&both) == 0) return -1;
  if (sched_setaffinity(0, sizeof(both), &both) != 0) return -2;
  return node;
}

extern "C" int pdb_bind_numa_for_pci(const char* pci_bus_id) {
  if (!pci_bus_id) return -1;
  std::string id(pci_bus_id);
  for (auto& c : id) c = static_cast<char>(std::tolower(static_cast<unsigned char>(c)));
  // cudaDeviceGetPCIBusId may give an 8-hex-digit domain ("00000000:1b:00.0"); sysfs uses 4.
  size_t colon = id.find(':');
  if (colon != std::string::npos && colon > 4) id = id.substr(colon - 4);
  std::string node_s;
  if (!read_file("/sys/bus/pci/devices/" + id + "/numa_node", &node_s)) return -1;
  return pdb_bind_numa_node(std::atoi(node_s.c_str()));
}

extern "C" int pdb_affinity_count(void) {
  cpu_set_t have;
  if (sched_getaffinity(0, sizeof(have), &have) != 0) return -1;
  return CPU_COUNT(&have);
}
