ssh test@127.0.0.1 "kill -9 3327"
ssh test@127.0.0.1 "kill -9 3328"
kill -9 2946
