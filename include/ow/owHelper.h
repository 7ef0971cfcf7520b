// owHelper.h — the part of the reference's owHelper (/root/reference/src/owHelper.h:15-45) that sits on
// the PCISPH path: configuration/*.txt loader, buffers/*.txt writer/reader and the stage stopwatch.
// The procedural worm generator (owHelper.cpp:104-1429) is out of scope.
#ifndef OW_HELPER_H
#define OW_HELPER_H

#include <time.h>

#include <string>

#include "owOpenCLConstant.h"

class owHelper {
 public:
  owHelper(void);
  ~owHelper(void);
  // Directory + file names are settable (the reference hard-codes ./configuration/position.txt etc.,
  // owHelper.cpp:1436,1466,1495,1513); defaults are the reference's.
  static void setConfigurationFiles(const char *dir, const char *position = "position.txt",
                                    const char *velocity = "velocity.txt",
                                    const char *connections = "elasticconnections.txt",
                                    const char *membranes = "membranes.txt",
                                    const char *particleMembranesList = "particlemembraneslist.txt");
  static void setBuffersDir(const char *dir);
  static void preLoadConfiguration();  // owHelper.cpp:1431-1458: counts rows, sets PARTICLE_COUNT(_RoundedUp)
  static void loadConfiguration(float *position, float *velocity, float *&elasticConnectionsData_cpp, int &numOfLiquidP,
                                int &numOfElasticP, int &numOfBoundaryP, int &numOfElasticConnections);  // :1460-1545
  // Extension: membranes for a file-loaded scene (the reference only gets them from its generator).
  // Returns false if the two files are absent. Allocates with new[].
  static bool loadMembranes(int *&membraneData_cpp, int &numOfMembranes, int *&particleMembranesList_cpp, int numOfElasticP);
  static void loadConfigurationToFile(float *position, float *connections = 0, int *membranes = 0,
                                      bool firstIteration = true);  // owHelper.cpp:1640-1672
  // Additions: the same recording off the caller's thread. The frame is copied (positions only: connections and
  // membranes belong to the first frame and must stay alive until flushRecording()), a writer thread formats and
  // appends it in order; flushRecording() blocks until everything queued is on disk.
  static void loadConfigurationToFileAsync(const float *position, float *connections = 0, int *membranes = 0,
                                           bool firstIteration = true);
  static void flushRecording();
  static void loadConfigurationFromFile(float *&position, float *&connections, int *&membranes,
                                        int iteration = 0);  // owHelper.cpp:1674-1740
  static void log_bufferf(const float *buffer, const int element_size, const int global_size, const char *fileName);
  static void log_bufferi(const int *buffer, const int element_size, const int global_size, const char *fileName);
  void watch_report(const char *str);  // owHelper.cpp:1806-1841
  double get_elapsedTime() { return elapsedTime; }
  void refreshTime();
  static bool quiet;  // suppress the per-stage printf lines (bench / tests)

 private:
  double elapsedTime;
  timespec t0, t1, t2;
};
#endif  // OW_HELPER_H
