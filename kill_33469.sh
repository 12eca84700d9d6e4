ssh test@127.0.0.1 "kill -9 9401"
ssh test@127.0.0.1 "kill -9 9402"
kill -9 8873
