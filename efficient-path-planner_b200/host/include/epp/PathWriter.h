// PathWriter.h — the reference's debug dumps (include/PathWriter.h:9-51, src/PathWriter.cpp:7-113) kept byte-compatible
// so that the plotting scripts written against `path_segments/` keep working:
//   path_<k>.txt, checkpoints.txt : one "x y z " line per point (stream default: %g, 6 significant digits)
//   gates.txt, obstacles.txt      : appended "id: <i> info: <row>" lines, the row in Eigen's default stream format
//                                   (every coefficient right-aligned to the widest, one space between)
// As in the reference an existing folder loses its regular files on construction.
#pragma once

#include <string>
#include <vector>

#include "LinAlg.h"

class PathWriter {
public:
    explicit PathWriter(const std::string &folderPath);
    void writePath(const std::vector<epp::Vector3d> &path);
    void updateGatePos(const int gateId, const epp::VectorXd &gateInfo);
    void updateObstaclePos(const int obstacleId, const epp::VectorXd &pose);
    void writeCheckpoints(const std::vector<epp::Vector3d> &checkpoints);

private:
    void dumpPoints(const std::string &name, const std::vector<epp::Vector3d> &points) const;
    void appendRow(const std::string &name, int id, const epp::VectorXd &row) const;
    int write_count = 0;
    const std::string folderPath;
};
