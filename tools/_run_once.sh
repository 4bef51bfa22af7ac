TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
COMMON="--synthetic --epochs 1 --max_steps 8"
rm -rf /tmp/ck && mkdir -p /tmp/ck
timeout 120 python main-single.py $COMMON --checkpoint_dir /tmp/ck/single 2>&1 | grep -a "validation\] Epoch\|Error\|error\|Traceback" | tail -3
timeout 120 $TR --master-port 29702 main-fsdp.py $COMMON --no_generate --checkpoint_dir /tmp/ck/fsdp 2>&1 | grep -a "validation\] Epoch\|Error\|error\|Traceback" | tail -3
timeout 120 $TR --master-port 29703 main-fsdp.py $COMMON --no_generate --reshard_after_forward --checkpoint_dir /tmp/ck/fsdp_rs 2>&1 | grep -a "validation\] Epoch\|Error\|error\|Traceback" | tail -3
timeout 120 $TR --master-port 29704 main-pipe.py $COMMON --no_generate --checkpoint_dir /tmp/ck/pipe 2>&1 | grep -a "validation\] Epoch\|Error\|error\|Traceback" | tail -3
timeout 120 $TR --master-port 29705 main-ddp.py $COMMON --no_generate --checkpoint_dir /tmp/ck/ddp 2>&1 | grep -a "validation\] Epoch\|Error\|error\|Traceback" | tail -3
ls /tmp/ck/*/ | head -20
