ssh test@127.0.0.1 "kill -9 8925"
ssh test@127.0.0.1 "kill -9 8926"
ssh test@127.0.0.1 "kill -9 8927"
kill -9 8873
