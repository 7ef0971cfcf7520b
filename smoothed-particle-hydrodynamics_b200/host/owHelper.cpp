// owHelper.cpp — configuration/*.txt loader, buffers/*.txt writer/reader, stage stopwatch.
// Formats and quirks follow /root/reference/src/owHelper.cpp:1431-1545 (loader), :1640-1740
// (record / playback), :44-57 + :1806-1841 (stopwatch). Paths are settable; defaults are the reference's.
#include "ow/owHelper.h"

#include <stdio.h>
#include <stdlib.h>

#include <fstream>
#include <iostream>
#include <stdexcept>
#include <vector>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>

#include "ow/owOpenCLSolver.h"

using namespace std;

bool owHelper::quiet = false;
static string g_cfg_dir = "./configuration", g_pos = "position.txt", g_vel = "velocity.txt",
              g_con = "elasticconnections.txt", g_mem = "membranes.txt", g_pml = "particlemembraneslist.txt";
static string g_buf_dir = "./buffers";

struct pos_rec { float x, y, z, p_type; };
static vector<pos_rec> f_data;

owHelper::owHelper(void) { refreshTime(); }
owHelper::~owHelper(void) {}

void owHelper::setConfigurationFiles(const char *dir, const char *position, const char *velocity,
                                     const char *connections, const char *membranes, const char *pml) {
  g_cfg_dir = dir; g_pos = position; g_vel = velocity; g_con = connections; g_mem = membranes; g_pml = pml;
}
void owHelper::setBuffersDir(const char *dir) { g_buf_dir = dir; }

void owHelper::refreshTime() {
  clock_gettime(CLOCK_MONOTONIC_RAW, &t1);
  t0 = t1;
}

// Prints `str` with the ms since the previous call and refreshes the elapsed time since refreshTime().
void owHelper::watch_report(const char *str) {
  clock_gettime(CLOCK_MONOTONIC_RAW, &t2);
  double ms = (double)(t2.tv_sec - t1.tv_sec) * 1000.0 + (double)(t2.tv_nsec - t1.tv_nsec) / 1000000.0;
  if (!quiet) printf(str, ms);
  t1 = t2;
  elapsedTime = (double)(t2.tv_sec - t0.tv_sec) * 1000.0 + (double)(t2.tv_nsec - t0.tv_nsec) / 1000000.0;
}

void owHelper::log_bufferf(const float *buffer, const int element_size, const int global_size, const char *fileName) {
  ofstream out(fileName);
  for (int i = 0; i < global_size; i++)
    for (int j = 0; j < element_size; j++) out << buffer[i * element_size + j] << (j < element_size - 1 ? "\t" : "\n");
}
void owHelper::log_bufferi(const int *buffer, const int element_size, const int global_size, const char *fileName) {
  ofstream out(fileName);
  for (int i = 0; i < global_size; i++)
    for (int j = 0; j < element_size; j++) out << buffer[i * element_size + j] << (j < element_size - 1 ? "\t" : "\n");
}

// Count parseable rows of position.txt: a row counts while its 4th field parses to a value >= 0.
void owHelper::preLoadConfiguration() {
  try {
    PARTICLE_COUNT = 0;
    ifstream positionFile((g_cfg_dir + "/" + g_pos).c_str());
    float x, y, z, p_type;
    if (!positionFile.is_open()) throw std::runtime_error("Could not open file " + g_pos);
    while (positionFile.good()) {
      p_type = -1.1f;
      positionFile >> x >> y >> z >> p_type;
      if (p_type >= 0) PARTICLE_COUNT++;
      else break;
    }
    PARTICLE_COUNT_RoundedUp = (((PARTICLE_COUNT - 1) / local_NDRange_size) + 1) * local_NDRange_size;
    if (!quiet)
      printf("\nConfiguration we are going to load contains %d particles. Now plan to allocate memory for them.\n", PARTICLE_COUNT);
  } catch (std::exception &e) {
    std::cout << "ERROR: " << e.what() << std::endl;
    exit(-1);
  }
}

void owHelper::loadConfiguration(float *position_cpp, float *velocity_cpp, float *&elasticConnections, int &nLiquid,
                                 int &nElastic, int &nBoundary, int &nConnections) {
  try {
    ifstream positionFile((g_cfg_dir + "/" + g_pos).c_str());
    int i = 0;
    float x, y, z, p_type;
    if (!positionFile.is_open()) throw std::runtime_error("Could not open file " + g_pos);
    while (positionFile.good() && i < PARTICLE_COUNT) {
      positionFile >> x >> y >> z >> p_type;
      position_cpp[4 * i + 0] = x; position_cpp[4 * i + 1] = y; position_cpp[4 * i + 2] = z; position_cpp[4 * i + 3] = p_type;
      switch ((int)p_type) {
        case LIQUID_PARTICLE: nLiquid++; break;
        case ELASTIC_PARTICLE: nElastic++; break;
        case BOUNDARY_PARTICLE: nBoundary++; break;
      }
      i++;
    }
    positionFile.close();
    ifstream velocityFile((g_cfg_dir + "/" + g_vel).c_str());
    if (!velocityFile.is_open()) throw std::runtime_error("Could not open file " + g_vel);
    i = 0;
    while (velocityFile.good() && i < PARTICLE_COUNT) {
      velocityFile >> x >> y >> z >> p_type;
      velocity_cpp[4 * i + 0] = x; velocity_cpp[4 * i + 1] = y; velocity_cpp[4 * i + 2] = z; velocity_cpp[4 * i + 3] = p_type;
      i++;
    }
    velocityFile.close();
    if (nElastic != 0) {  // owHelper.cpp:1511-1540: rows "jd rij0 val1 val2", rows with jd < -1 skipped
      ifstream conFile((g_cfg_dir + "/" + g_con).c_str());
      elasticConnections = new float[4 * nElastic * MAX_NEIGHBOR_COUNT];
      for (int k = 0; k < nElastic * MAX_NEIGHBOR_COUNT; k++) {  // absent file => no springs (reference leaves it uninitialised)
        elasticConnections[4 * k + 0] = NO_PARTICLE_ID; elasticConnections[4 * k + 1] = 0;
        elasticConnections[4 * k + 2] = 0; elasticConnections[4 * k + 3] = 0;
      }
      i = 0;
      float jd, rij0, val1, val2;
      if (conFile.is_open()) {
        while (conFile.good() && i < nElastic * MAX_NEIGHBOR_COUNT) {
          jd = -10;
          conFile >> jd >> rij0 >> val1 >> val2;
          if (jd >= -1) {
            elasticConnections[4 * i + 0] = jd; elasticConnections[4 * i + 1] = rij0;
            elasticConnections[4 * i + 2] = val1; elasticConnections[4 * i + 3] = val2;
            if (jd >= 0) nConnections++;
            i++;
          }
        }
      }
    }
  } catch (std::exception &e) {
    std::cout << "ERROR: " << e.what() << std::endl;
    exit(-1);
  }
}

// membranes.txt: first line M, then M rows "i j k"; particlemembraneslist.txt: Ne rows of 7 ints.
bool owHelper::loadMembranes(int *&membraneData_cpp, int &nMembranes, int *&pml, int nElastic) {
  ifstream m((g_cfg_dir + "/" + g_mem).c_str()), l((g_cfg_dir + "/" + g_pml).c_str());
  if (!m.is_open() || !l.is_open() || nElastic <= 0) return false;
  m >> nMembranes;
  if (nMembranes <= 0) return false;
  membraneData_cpp = new int[3 * nMembranes];
  for (int i = 0; i < 3 * nMembranes; i++) m >> membraneData_cpp[i];
  pml = new int[MAX_MEMBRANES_INCLUDING_SAME_PARTICLE * nElastic];
  for (int i = 0; i < MAX_MEMBRANES_INCLUDING_SAME_PARTICLE * nElastic; i++) { pml[i] = -1; l >> pml[i]; }
  return true;
}

// buffers/position_buffer.txt: line 1 Ne, line 2 Nl, then one row "x\ty\tz\ttype" per NON-boundary particle per
// saved frame; connection_buffer.txt: 32*Ne rows of 4 floats; membranes_buffer.txt: count, then rows.
// membranes_buffer.txt keeps the reference's row width of FOUR ints (owHelper.cpp:1664-1665 writes, :1724-1727
// reads four per row) so that files travel between the two programs; the reference fills the rows from a 3-stride
// array with index 4*i (rows shifted, the last quarter read out of bounds): here a row is the triangle's 3 indices
// plus a 0 filler. The reader below accepts 3 or 4 columns.
void owHelper::loadConfigurationToFile(float *position, float *connections, int *membranes, bool firstIteration) {
  try {
    ofstream positionFile;
    string pf = g_buf_dir + "/position_buffer.txt";
    if (firstIteration) {
      positionFile.open(pf.c_str(), std::ofstream::trunc);
      positionFile << numOfElasticP << "\n";
      positionFile << numOfLiquidP << "\n";
    } else {
      positionFile.open(pf.c_str(), std::ofstream::app);
    }
    if (!positionFile.is_open()) throw std::runtime_error("Could not open file " + pf);
    for (int i = 0; i < PARTICLE_COUNT; i++)
      if ((int)position[4 * i + 3] != BOUNDARY_PARTICLE)
        positionFile << position[i * 4 + 0] << "\t" << position[i * 4 + 1] << "\t" << position[i * 4 + 2] << "\t"
                     << position[i * 4 + 3] << "\n";
    positionFile.close();
    if (firstIteration) {
      ofstream connectionFile((g_buf_dir + "/connection_buffer.txt").c_str(), std::ofstream::trunc);
      int con_num = MAX_NEIGHBOR_COUNT * numOfElasticP;
      if (connections)
        for (int i = 0; i < con_num; i++)
          connectionFile << connections[4 * i + 0] << "\t" << connections[4 * i + 1] << "\t" << connections[4 * i + 2]
                         << "\t" << connections[4 * i + 3] << "\n";
      connectionFile.close();
      ofstream membranesFile((g_buf_dir + "/membranes_buffer.txt").c_str(), std::ofstream::trunc);
      membranesFile << numOfMembranes << "\n";
      if (membranes)
        for (int i = 0; i < numOfMembranes; i++)
          membranesFile << membranes[3 * i + 0] << "\t" << membranes[3 * i + 1] << "\t" << membranes[3 * i + 2] << "\t" << 0 << "\n";
      membranesFile.close();
    }
  } catch (std::exception &e) {
    std::cout << "ERROR: " << e.what() << std::endl;
    exit(-1);
  }
}

// ---- asynchronous recording: one writer thread, frames in call order ----
namespace {
struct RecJob { std::vector<float> position; float *connections; int *membranes; bool first; };
// heap objects that are never destroyed: the detached writer thread waits on them for the life of the process, and
// destroying a condition variable with a waiter at exit (static destructors) blocks forever in pthread_cond_destroy
std::mutex &g_rec_mu = *new std::mutex;
std::condition_variable &g_rec_cv = *new std::condition_variable, &g_rec_idle = *new std::condition_variable;
std::deque<RecJob> &g_rec_q = *new std::deque<RecJob>;
bool g_rec_busy = false, g_rec_started = false;
void rec_worker() {
  for (;;) {
    RecJob job;
    {
      std::unique_lock<std::mutex> lk(g_rec_mu);
      g_rec_cv.wait(lk, [] { return !g_rec_q.empty(); });
      job = std::move(g_rec_q.front());
      g_rec_q.pop_front();
      g_rec_busy = true;
    }
    owHelper::loadConfigurationToFile(job.position.data(), job.connections, job.membranes, job.first);
    {
      std::lock_guard<std::mutex> lk(g_rec_mu);
      g_rec_busy = false;
    }
    g_rec_idle.notify_all();
  }
}
}  // namespace

void owHelper::loadConfigurationToFileAsync(const float *position, float *connections, int *membranes, bool firstIteration) {
  RecJob job;
  job.position.assign(position, position + 4 * (size_t)PARTICLE_COUNT);
  job.connections = connections; job.membranes = membranes; job.first = firstIteration;
  {
    std::lock_guard<std::mutex> lk(g_rec_mu);
    if (!g_rec_started) { std::thread(rec_worker).detach(); g_rec_started = true; }
    g_rec_q.push_back(std::move(job));
  }
  g_rec_cv.notify_one();
}

void owHelper::flushRecording() {
  std::unique_lock<std::mutex> lk(g_rec_mu);
  g_rec_idle.wait(lk, [] { return g_rec_q.empty() && !g_rec_busy; });
}

// Playback reader (owHelper.cpp:1674-1740): frame `iteration` of position_buffer.txt; connections and
// membranes on iteration 0. Sets PARTICLE_COUNT = Ne + Nl and iterationCount = number of frames.
void owHelper::loadConfigurationFromFile(float *&position, float *&connections, int *&membranes, int iteration) {
  flushRecording();
  try {
    if (iteration == 0) {
      f_data.clear();
      ifstream positionFile((g_buf_dir + "/position_buffer.txt").c_str());
      if (!positionFile.is_open()) throw std::runtime_error("Could not open file position_buffer.txt");
      float x, y, z, p_type;
      positionFile >> numOfElasticP;
      positionFile >> numOfLiquidP;
      while (positionFile >> x >> y >> z >> p_type) {
        pos_rec p = {x, y, z, p_type};
        f_data.push_back(p);
      }
      PARTICLE_COUNT = numOfElasticP + numOfLiquidP;
      iterationCount = PARTICLE_COUNT > 0 ? (int)(f_data.size() / PARTICLE_COUNT) : 0;
      position = new float[4 * PARTICLE_COUNT];
    }
    if ((size_t)(iteration + 1) * PARTICLE_COUNT > f_data.size()) throw std::runtime_error("no such frame");
    for (int i = 0; i < PARTICLE_COUNT; i++) {
      const pos_rec &p = f_data[i + (size_t)PARTICLE_COUNT * iteration];
      position[i * 4 + 0] = p.x; position[i * 4 + 1] = p.y; position[i * 4 + 2] = p.z; position[i * 4 + 3] = p.p_type;
    }
    if (iteration == 0) {
      ifstream connectionFile((g_buf_dir + "/connection_buffer.txt").c_str());
      connections = new float[MAX_NEIGHBOR_COUNT * numOfElasticP * 4 + 4];
      int i = 0;
      float jd, rij0, val1, val2;
      while (connectionFile.is_open() && i < MAX_NEIGHBOR_COUNT * numOfElasticP && (connectionFile >> jd >> rij0 >> val1 >> val2)) {
        connections[4 * i + 0] = jd; connections[4 * i + 1] = rij0; connections[4 * i + 2] = val1; connections[4 * i + 3] = val2;
        i++;
      }
      ifstream membranesFile((g_buf_dir + "/membranes_buffer.txt").c_str());
      int m_count = 0;
      if (membranesFile.is_open()) membranesFile >> m_count;
      numOfMembranes = m_count;
      membranes = new int[3 * m_count + 3];
      std::vector<int> tok;
      for (int v; membranesFile >> v;) tok.push_back(v);
      const int cols = (m_count > 0 && tok.size() >= (size_t)4 * m_count) ? 4 : 3;  // reference row width, or the 3-column files of round 1
      for (i = 0; i < m_count && (size_t)(cols * i + 2) < tok.size(); i++) {
        membranes[3 * i] = tok[cols * i]; membranes[3 * i + 1] = tok[cols * i + 1]; membranes[3 * i + 2] = tok[cols * i + 2];
      }
    }
  } catch (std::exception &e) {
    std::cout << "ERROR: " << e.what() << std::endl;
    exit(-1);
  }
}
