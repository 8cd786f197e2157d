g++ -O2 -std=c++17 -pthread -o /tmp/ph tools/uring_setup_probe.cc || exit 1
echo "parallel"; for i in $(seq 1 25); do /tmp/ph 16 0; done | sort | uniq -c | sort -k4 -n | tail -8
echo "serial"; for i in $(seq 1 25); do /tmp/ph 16 1; done | sort -k3 -n | tail -6
