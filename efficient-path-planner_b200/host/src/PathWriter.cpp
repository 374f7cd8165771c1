// PathWriter.cpp — see PathWriter.h.  C stdio throughout: "%g" is what operator<<(double) prints by default.
#include "epp/PathWriter.h"

#include <algorithm>
#include <cstdio>
#include <filesystem>

namespace fs = std::filesystem;

namespace {
std::string cell(double v)
{
    char buf[40];
    std::snprintf(buf, sizeof buf, "%g", v);
    return buf;
}
} // namespace

PathWriter::PathWriter(const std::string &folderPath) : folderPath(folderPath)
{
    if (fs::create_directories(folderPath)) {
        std::printf("Folder created: %s\n", folderPath.c_str());
        return;
    }
    std::vector<fs::path> stale;
    for (const fs::directory_entry &e : fs::directory_iterator(folderPath))
        if (e.is_regular_file() && !e.is_symlink()) stale.push_back(e.path());
    for (const fs::path &p : stale) fs::remove(p);
}

void PathWriter::dumpPoints(const std::string &name, const std::vector<epp::Vector3d> &points) const
{
    std::FILE *f = std::fopen((folderPath + "/" + name).c_str(), "w");
    if (!f) {
        std::fprintf(stderr, "Failed to open file for writing: %s\n", name.c_str());
        return;
    }
    for (const epp::Vector3d &p : points) std::fprintf(f, "%g %g %g \n", p(0), p(1), p(2));
    std::fclose(f);
}

void PathWriter::appendRow(const std::string &name, int id, const epp::VectorXd &row) const
{
    std::FILE *f = std::fopen((folderPath + "/" + name).c_str(), "a");
    if (!f) {
        std::fprintf(stderr, "Failed to open file for writing: %s\n", name.c_str());
        return;
    }
    std::vector<std::string> cells;
    std::size_t width = 0;
    for (int i = 0; i < (int)row.size(); ++i) {
        cells.push_back(cell(row(i)));
        width = std::max(width, cells.back().size());
    }
    std::fprintf(f, "id: %d info: ", id);
    for (std::size_t i = 0; i < cells.size(); ++i) std::fprintf(f, "%s%*s", i ? " " : "", (int)width, cells[i].c_str());
    std::fputc('\n', f);
    std::fclose(f);
}

void PathWriter::writePath(const std::vector<epp::Vector3d> &path)
{
    dumpPoints("path_" + std::to_string(write_count++) + ".txt", path);
}

void PathWriter::updateGatePos(const int gateId, const epp::VectorXd &gateInfo) { appendRow("gates.txt", gateId, gateInfo); }

void PathWriter::updateObstaclePos(const int obstacleId, const epp::VectorXd &pose)
{
    appendRow("obstacles.txt", obstacleId, pose);
}

void PathWriter::writeCheckpoints(const std::vector<epp::Vector3d> &checkpoints) { dumpPoints("checkpoints.txt", checkpoints); }
