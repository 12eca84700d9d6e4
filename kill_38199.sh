ssh test@127.0.0.1 "kill -9 3748"
ssh test@127.0.0.1 "kill -9 3749"
kill -9 2946
