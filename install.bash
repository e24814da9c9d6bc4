# Same interface as the reference's install.bash (/root/reference/install.bash:4-20):
#     source install.bash [STATE_SIZE CONTROL_SIZE KNOT_POINTS]
# builds build/gpu_library*.so with cmake and puts build/ on PYTHONPATH.  (No git submodules here: the pybind11 headers
# come from the Python environment.)
mkdir -p build
cd build
if [ -z $1 ]; then
        echo "Using default values of STATE_SIZE, CONTROL_SIZE, KNOT_POINTS"
        cmake .. -DS=14 -DC=7 -DK=50
elif [ -z $2 ]; then
        echo "Not enough arguments, please specify STATE_SIZE, CONTROL_SIZE, KNOT_POINTS"
        cd ..
        return
elif [ -z $3 ]; then
        echo "Not enough arguments, please specify STATE_SIZE, CONTROL_SIZE, KNOT_POINTS"
        cd ..
        return
else
        cmake .. -DS=$1 -DC=$2 -DK=$3
fi
make
export PYTHONPATH=$PWD:$PYTHONPATH
cd ..
